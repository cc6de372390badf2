{-# LANGUAGE GADTs               #-}
{-# LANGUAGE ScopedTypeVariables #-}
{-# LANGUAGE TypeApplications    #-}
-- |
-- Module      : Data.Array.Accelerate.Numeric.LinearAlgebra.LLVM.PTX.B200.Level1
-- Description : NEW ForeignAcc hooks for the Level-1 reductions of BLAS/Level1.hs
--
-- SOURCE ONLY (no GHC in the build image; never compiled).  The reference evaluates
-- sdot / asum / amax / amin as Accelerate folds (BLAS/Level1.hs:44-62, 100-143); each becomes
--
-- > asum = foreignAcc (B200.asum nR) $ \xs -> sum (map mag xs)          -- fallback kept
--
-- so that the llvm-ptx backend runs the one-pass kernel of libaccblas_b200.so and every other
-- backend keeps the fold.  The alpha-on-device variants of gemm / gemv live in Hooks.hs.
--
module Data.Array.Accelerate.Numeric.LinearAlgebra.LLVM.PTX.B200.Level1 ( asum, amax, amin ) where

import Data.Array.Accelerate.Numeric.LinearAlgebra.Type
import Data.Array.Accelerate.Numeric.LinearAlgebra.LLVM.PTX.Base            ( withArray )
import Data.Array.Accelerate.Numeric.LinearAlgebra.LLVM.PTX.B200.FFI
import Data.Array.Accelerate.Numeric.LinearAlgebra.LLVM.PTX.B200.Hooks      ( tag, raw, streamPtr )

import Data.Array.Accelerate.LLVM.PTX.Foreign
import Data.Array.Accelerate.Representation.Array
import Data.Array.Accelerate.Representation.Shape
import Data.Array.Accelerate.Sugar.Elt

import Control.Monad.Reader
import Foreign.Ptr                                                          ( Ptr )

type Reduce1 = Ptr () -> CInt -> Int64 -> Ptr () -> Ptr () -> IO Status

-- | One-vector reduction into a device Scalar of element type @r@ (NumericBaseT e for asum,
-- Int for amax / amin -- the reference's `Scalar Int`, a 64-bit index).
reduce1 :: String -> Reduce1 -> NumericR s e -> TypeR r -> ForeignAcc (Vector e -> Scalar r)
reduce1 name c_fn nR rR = ForeignAcc name $ \xs -> do
  let ((), n) = shape xs
  future <- new
  stream <- asks ptxStream
  out    <- allocateRemote (ArrayR dim0 rR) ()
  ()     <- liftPar $
    withArray nR xs stream $ \pX ->
      withScalarArray rR out stream $ \pO -> liftIO $          -- raw device pointer of the result
        checkStatus name =<< c_fn (streamPtr stream) (dtypeCode (tag nR)) (fromIntegral n) (raw pX) (raw pO)
  put future out
  return future

asum :: NumericR s e -> TypeR (NumericBaseT e) -> ForeignAcc (Vector e -> Scalar (NumericBaseT e))
asum = reduce1 "ptx.asum" c_accblas_asum

-- | Index of the element of largest magnitude; on ties the LAST one, as the reference's
-- `cmp ix iy = snd ix > snd iy ? ix : iy` (BLAS/Level1.hs:124) returns.
amax, amin :: NumericR s e -> ForeignAcc (Vector e -> Scalar Int)
amax nR = reduce1 "ptx.amax" c_accblas_amax nR (eltR @Int)
amin nR = reduce1 "ptx.amin" c_accblas_amin nR (eltR @Int)
