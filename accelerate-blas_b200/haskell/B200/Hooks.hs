{-# LANGUAGE GADTs               #-}
{-# LANGUAGE ScopedTypeVariables #-}
{-# LANGUAGE TypeApplications    #-}
-- |
-- Module      : Data.Array.Accelerate.Numeric.LinearAlgebra.LLVM.PTX.B200.Hooks
-- Description : the three ForeignAcc hooks of the llvm-ptx backend, calling libaccblas_b200.so
--
-- SOURCE ONLY (no GHC in the build image; never compiled).  Drop-in for
--   LLVM/PTX/Level3.hs  gemm   (ForeignAcc "ptx.gemm")
--   LLVM/PTX/Level2.hs  gemv   (ForeignAcc "ptx.gemv")
-- plus the NEW hook  dot  (ForeignAcc "ptx.dot") for BLAS/Level1.hs.
--
-- What stays as in the reference: shape arithmetic, `allocateRemote` of the (uninitialised) result,
-- `indexRemote` of alpha, `withArray` (raw device pointers + the post-call waypoint event that
-- keeps the buffers alive, LLVM/PTX/Base.hs:94-98), the `Future` protocol.
-- What goes away: `withBLAS` and the cuBLAS handle cache (LLVM/PTX/Context.hs), `withV2`
-- (alpha is passed by plain host pointer for every dtype), the operand swap for column-major
-- cuBLAS, and the gemm detour for complex-H gemv (LLVM/PTX/Level2.hs:52-69).
-- What is new: the backend's stream is passed down (the reference never called setStream).
--
module Data.Array.Accelerate.Numeric.LinearAlgebra.LLVM.PTX.B200.Hooks ( gemm, gemmAlphaDev, gemv, dot, tag, raw, streamPtr ) where

import Data.Array.Accelerate.Numeric.LinearAlgebra.Type
import Data.Array.Accelerate.Numeric.LinearAlgebra.LLVM.PTX.Base            ( withArray, withV2 )
import Data.Array.Accelerate.Numeric.LinearAlgebra.LLVM.PTX.B200.FFI

import Data.Array.Accelerate.Array.Data                                     ( ArrayData )
import Data.Array.Accelerate.Data.Complex
import Data.Array.Accelerate.LLVM.PTX.Foreign
import Data.Array.Accelerate.Representation.Array
import Data.Array.Accelerate.Representation.Shape
import Data.Array.Accelerate.Sugar.Elt
import Data.Array.Accelerate.Type

import Control.Monad.Reader
import Foreign.Marshal.Utils                                                ( with )
import Foreign.Ptr                                                          ( Ptr, castPtr )
import qualified Foreign.CUDA.Ptr                                           as CUDA
import qualified Foreign.CUDA.Driver.Stream                                 as CUDA

tag :: NumericR s e -> DType
tag NumericRfloat32   = F32
tag NumericRfloat64   = F64
tag NumericRcomplex32 = C32
tag NumericRcomplex64 = C64

raw :: CUDA.DevicePtr a -> Ptr ()
raw = castPtr . CUDA.useDevicePtr

streamPtr :: Stream -> Ptr ()            -- CUstream is a pointer newtype in the `cuda` package
streamPtr = castPtr . CUDA.useStream . unsafeGetValue

-- | C = alpha * op(A) * op(B)
gemm :: NumericR s e -> Transpose -> Transpose
     -> ForeignAcc (((((), Scalar e), Matrix e), Matrix e) -> Matrix e)
gemm nR opA opB = ForeignAcc "ptx.gemm" $ \((((), alpha), matA), matB) -> do
  let (((), rowsA), colsA) = shape matA
      (((), rowsB), colsB) = shape matB
      m = if opA == N then rowsA else colsA
      n = if opB == N then colsB else rowsB
      eR = eltOf nR
  future <- new
  stream <- asks ptxStream
  matC   <- allocateRemote (ArrayR dim2 eR) (((), m), n)
  alpha' <- indexRemote eR alpha 0                  -- host copy of alpha, as in the reference
  ()     <- liftPar $
    withArray nR matA stream   $ \pA ->
     withArray nR matB stream  $ \pB ->
      withArray nR matC stream $ \pC -> liftIO $
        withScalar nR alpha' $ \pAlpha ->
          checkStatus "accblas_gemm" =<<
            c_accblas_gemm (streamPtr stream) (dtypeCode (tag nR)) (opCode opA) (opCode opB)
                           (fromIntegral rowsA) (fromIntegral colsA) (fromIntegral rowsB) (fromIntegral colsB)
                           pAlpha (raw pA) (raw pB) (raw pC)
  put future matC
  return future

-- | The same hook without the blocking `indexRemote`: alpha stays the device Scalar that
-- BLAS/Level3.hs:64 builds (`unit alpha`) and the kernel reads it (accblas_gemm_alpha_dev).  This removes
-- the one host synchronisation the reference pays per product (LLVM/PTX/Level3.hs:81).
gemmAlphaDev :: NumericR s e -> Transpose -> Transpose
             -> ForeignAcc (((((), Scalar e), Matrix e), Matrix e) -> Matrix e)
gemmAlphaDev nR opA opB = ForeignAcc "ptx.gemm" $ \((((), alpha), matA), matB) -> do
  let (((), rowsA), colsA) = shape matA
      (((), rowsB), colsB) = shape matB
      m = if opA == N then rowsA else colsA
      n = if opB == N then colsB else rowsB
  future <- new
  stream <- asks ptxStream
  matC   <- allocateRemote (ArrayR dim2 (eltOf nR)) (((), m), n)
  ()     <- liftPar $
    withArray nR alpha stream    $ \pAlpha ->
     withArray nR matA stream    $ \pA ->
      withArray nR matB stream   $ \pB ->
       withArray nR matC stream  $ \pC -> liftIO $
          checkStatus "accblas_gemm_alpha_dev" =<<
            c_accblas_gemm_alpha_dev (streamPtr stream) (dtypeCode (tag nR)) (opCode opA) (opCode opB)
                                     (fromIntegral rowsA) (fromIntegral colsA) (fromIntegral rowsB) (fromIntegral colsB)
                                     (raw pAlpha) (raw pA) (raw pB) (raw pC)
  put future matC
  return future

-- | y = alpha * op(A) * x   (all four dtypes, all three ops: no gemm detour)
gemv :: NumericR s e -> Transpose
     -> ForeignAcc ((((), Scalar e), Matrix e), Vector e) -> Vector e)
gemv nR opA = ForeignAcc "ptx.gemv" $ \((((), alpha), matA), vecx) -> do
  let (((), rowsA), colsA) = shape matA
      ((), lenX)           = shape vecx
      lenY = if opA == N then rowsA else colsA
      eR   = eltOf nR
  future <- new
  stream <- asks ptxStream
  vecy   <- allocateRemote (ArrayR dim1 eR) ((), lenY)
  alpha' <- indexRemote eR alpha 0
  ()     <- liftPar $
    withArray nR matA stream   $ \pA ->
     withArray nR vecx stream  $ \pX ->
      withArray nR vecy stream $ \pY -> liftIO $
        withScalar nR alpha' $ \pAlpha ->
          checkStatus "accblas_gemv" =<<
            c_accblas_gemv (streamPtr stream) (dtypeCode (tag nR)) (opCode opA)
                           (fromIntegral rowsA) (fromIntegral colsA)
                           pAlpha (raw pA) (raw pX) (fromIntegral lenX) (raw pY)
  put future vecy
  return future

-- | NEW.  dotu (conj = False) / dotc (conj = True); result stays on the device as a Scalar.
dot :: NumericR s e -> Bool -> ForeignAcc ((((), Vector e), Vector e) -> Scalar e)
dot nR conj = ForeignAcc "ptx.dot" $ \(((), xs), ys) -> do
  let ((), nx) = shape xs
      ((), ny) = shape ys
      eR       = eltOf nR
  future <- new
  stream <- asks ptxStream
  out    <- allocateRemote (ArrayR dim0 eR) ()
  ()     <- liftPar $
    withArray nR xs stream   $ \pX ->
     withArray nR ys stream  $ \pY ->
      withArray nR out stream $ \pO -> liftIO $
        checkStatus "accblas_dot" =<<
          c_accblas_dot (streamPtr stream) (dtypeCode (tag nR)) (if conj then 1 else 0)
                        (fromIntegral (min nx ny))        -- zipWith intersects the extents
                        (raw pX) (raw pY) (raw pO)
  put future out
  return future

-- helpers -------------------------------------------------------------------------------------

eltOf :: NumericR s e -> TypeR e
eltOf NumericRfloat32   = eltR @Float
eltOf NumericRfloat64   = eltR @Double
eltOf NumericRcomplex32 = eltR @(Complex Float)
eltOf NumericRcomplex64 = eltR @(Complex Double)

-- | alpha by host pointer: `with` marshals a Storable scalar; a complex alpha is the two-element
-- Vec2 the reference already pins with withV2 (LLVM/PTX/Base.hs:50-75), reused unchanged.
withScalar :: NumericR s e -> e -> (Ptr () -> IO b) -> IO b
withScalar NumericRfloat32   a k = with a (k . castPtr)
withScalar NumericRfloat64   a k = with a (k . castPtr)
withScalar NumericRcomplex32 a k = withV2 NumericRcomplex32 a (k . castPtr)
withScalar NumericRcomplex64 a k = withV2 NumericRcomplex64 a (k . castPtr)
