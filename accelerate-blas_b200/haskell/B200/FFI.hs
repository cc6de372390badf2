{-# LANGUAGE ForeignFunctionInterface #-}
-- |
-- Module      : Data.Array.Accelerate.Numeric.LinearAlgebra.LLVM.PTX.B200.FFI
-- Description : raw bindings to libaccblas_b200.so (include/accblas_b200.h)
--
-- SOURCE ONLY: the build image has no GHC, so this module has never been compiled.  It shows the
-- exact binding a maintainer adds; the executable verification of the C ABI runs through the
-- ctypes mirror (accelerate-blas_b200/python/accblas_b200/_ffi.py).
--
-- cabal: add `extra-libraries: accblas_b200` to the llvm-ptx stanza and drop the `cublas`
-- dependency (accelerate-blas.cabal:74-84).
--
module Data.Array.Accelerate.Numeric.LinearAlgebra.LLVM.PTX.B200.FFI (

  DType(..), dtypeCode, opCode, Status, checkStatus,
  c_accblas_gemm, c_accblas_gemv, c_accblas_dot, c_accblas_dot_compensated, c_accblas_dot_static,
  c_accblas_gemm_alpha_dev, c_accblas_gemv_alpha_dev,
  c_accblas_sdot, c_accblas_asum, c_accblas_amax, c_accblas_amin,
  c_accblas_gemm_strided_batched, c_accblas_set_tuning,
  c_accblas_gemm_ex, c_accblas_trim_workspace, c_accblas_last_cuda_error_r,
  -- multi-GPU (one process per device)
  Comm, Plan, c_accblas_mg_unique_id, c_accblas_mg_init, c_accblas_mg_finalize, c_accblas_mg_row_panel,
  c_accblas_mg_gemm, c_accblas_mg_gemm_plan, c_accblas_mg_gemm_run, c_accblas_mg_gemm_plan_destroy,
  c_accblas_mg_gemv, c_accblas_mg_dot,

) where

import Data.Array.Accelerate.Numeric.LinearAlgebra.Type            ( Transpose(..) )

import Control.Exception                                            ( throwIO, ErrorCall(..) )
import Control.Monad                                                ( when )
import Data.Int                                                     ( Int64 )
import Foreign.C.String                                             ( CString, peekCString )
import Foreign.C.Types                                              ( CInt(..), CSize(..) )
import Foreign.Ptr                                                  ( Ptr )

-- | accblas_dtype_t; one constructor per NumericR constructor (Type.hs:32-36)
data DType = F32 | F64 | C32 | C64

dtypeCode :: DType -> CInt
dtypeCode F32 = 0
dtypeCode F64 = 1
dtypeCode C32 = 2
dtypeCode C64 = 3

-- | accblas_op_t (Type.hs:78-90)
opCode :: Transpose -> CInt
opCode N = 0
opCode T = 1
opCode H = 2

type Status = CInt

-- The CUstream the backend hands out (`asks ptxStream`) is passed as an opaque pointer; device
-- buffers are the raw addresses behind `DevicePtr` (useDevicePtr).
foreign import ccall unsafe "accblas_gemm"
  c_accblas_gemm :: Ptr () -> CInt -> CInt -> CInt          -- stream, dtype, opA, opB
                 -> Int64 -> Int64 -> Int64 -> Int64        -- rowsA colsA rowsB colsB
                 -> Ptr () -> Ptr () -> Ptr () -> Ptr ()    -- alpha(host) A B C
                 -> IO Status

foreign import ccall unsafe "accblas_gemv"
  c_accblas_gemv :: Ptr () -> CInt -> CInt                  -- stream, dtype, opA
                 -> Int64 -> Int64                          -- rowsA colsA
                 -> Ptr () -> Ptr () -> Ptr () -> Int64     -- alpha(host) A x |x|
                 -> Ptr () -> IO Status                     -- y

foreign import ccall unsafe "accblas_dot"
  c_accblas_dot  :: Ptr () -> CInt -> CInt -> Int64         -- stream, dtype, conj_x, n
                 -> Ptr () -> Ptr () -> Ptr ()              -- x y out(device)
                 -> IO Status

-- the same product with compensated arithmetic (Dot2 + the KBN combine of Numeric/Sum.hs:119-125)
foreign import ccall unsafe "accblas_dot_compensated"
  c_accblas_dot_compensated :: Ptr () -> CInt -> CInt -> Int64
                            -> Ptr () -> Ptr () -> Ptr ()
                            -> IO Status

-- accblas_dot with its partial sums in a fixed array inside the library: allocates nothing, so it cannot disturb
-- the backend's memory table (Data.Array.Accelerate.LLVM.PTX.Array.Remote).  One call at a time per device.
foreign import ccall unsafe "accblas_dot_static"
  c_accblas_dot_static :: Ptr () -> CInt -> CInt -> Int64
                       -> Ptr () -> Ptr () -> Ptr ()
                       -> IO Status

-- alpha as the device Scalar it already is (BLAS/Level3.hs:64): no indexRemote, no stream sync
foreign import ccall unsafe "accblas_gemm_alpha_dev"
  c_accblas_gemm_alpha_dev :: Ptr () -> CInt -> CInt -> CInt
                           -> Int64 -> Int64 -> Int64 -> Int64
                           -> Ptr () -> Ptr () -> Ptr () -> Ptr ()    -- alpha(DEVICE) A B C
                           -> IO Status

foreign import ccall unsafe "accblas_gemv_alpha_dev"
  c_accblas_gemv_alpha_dev :: Ptr () -> CInt -> CInt
                           -> Int64 -> Int64
                           -> Ptr () -> Ptr () -> Ptr () -> Int64     -- alpha(DEVICE) A x |x|
                           -> Ptr () -> IO Status

-- the remaining Level-1 reductions (BLAS/Level1.hs:44-62, 100-143); results in device memory
foreign import ccall unsafe "accblas_sdot"
  c_accblas_sdot :: Ptr () -> CInt -> Int64 -> Ptr ()       -- stream, dtype, n, z(host)
                 -> Ptr () -> Ptr () -> Ptr () -> IO Status -- x y out(device)

foreign import ccall unsafe "accblas_asum"
  c_accblas_asum :: Ptr () -> CInt -> Int64 -> Ptr () -> Ptr () -> IO Status   -- out: base type

foreign import ccall unsafe "accblas_amax"
  c_accblas_amax :: Ptr () -> CInt -> Int64 -> Ptr () -> Ptr () -> IO Status   -- out: Int (64-bit)

foreign import ccall unsafe "accblas_amin"
  c_accblas_amin :: Ptr () -> CInt -> Int64 -> Ptr () -> Ptr () -> IO Status

-- strided batch of equal-shaped products (no operator of the reference uses it yet; a future batched
-- `(<>)` over rank-3 arrays would): strides in elements, stride 0 = the same operand for every product
foreign import ccall unsafe "accblas_gemm_strided_batched"
  c_accblas_gemm_strided_batched :: Ptr () -> CInt -> CInt -> CInt
                                 -> Int64 -> Int64 -> Int64 -> Int64
                                 -> Ptr () -> Ptr () -> Int64 -> Ptr () -> Int64    -- alpha(host) A strideA B strideB
                                 -> Ptr () -> Int64 -> Int64 -> IO Status           -- C strideC batch

-- the extended product D = alpha * op(A) * op(B) + beta * Cin on windows (row pitches in elements); not used by the
-- reference's hooks (beta = 0, LLVM/PTX/Level3.hs:90) but by a fused `zipWith (+) c (a <> b)` and by the multi-GPU path
foreign import ccall unsafe "accblas_gemm_ex"
  c_accblas_gemm_ex :: Ptr () -> CInt -> CInt -> CInt
                    -> Int64 -> Int64 -> Int64 -> Int64
                    -> Ptr () -> Ptr () -> Int64 -> Ptr () -> Int64       -- alpha(host) A lda B ldb
                    -> Ptr () -> Ptr () -> Int64 -> Ptr () -> Int64       -- beta(host) Cin ldcin C ldc
                    -> IO Status

-- multi-GPU: one OS process (one PTX context) per device of an NVLink box.  The 128 bootstrap bytes of rank 0 travel
-- to the other ranks by whatever transport the host runtime has (a file, a socket, MPI).
type Comm = Ptr ()
type Plan = Ptr ()

foreign import ccall unsafe "accblas_mg_unique_id"
  c_accblas_mg_unique_id :: Ptr () -> IO Status                           -- 128 bytes out (rank 0)

foreign import ccall safe "accblas_mg_init"                               -- collective, blocks: `safe`
  c_accblas_mg_init :: Ptr Comm -> CInt -> CInt -> Ptr () -> IO Status    -- comm(out) rank world id

foreign import ccall safe "accblas_mg_finalize"
  c_accblas_mg_finalize :: Comm -> IO Status

foreign import ccall unsafe "accblas_mg_row_panel"
  c_accblas_mg_row_panel :: Int64 -> CInt -> CInt -> Ptr Int64 -> Ptr Int64 -> IO Status   -- m world rank start(out) rows(out)

-- C = alpha * op(A) * op(B): operands and result on `root`, rows of op(A) / C split over the ranks
foreign import ccall safe "accblas_mg_gemm"
  c_accblas_mg_gemm :: Comm -> Ptr () -> CInt -> CInt -> CInt
                    -> Int64 -> Int64 -> Int64 -> Int64
                    -> Ptr () -> Ptr () -> Ptr () -> Ptr () -> CInt       -- alpha(host) A B C root
                    -> IO Status

foreign import ccall safe "accblas_mg_gemm_plan"
  c_accblas_mg_gemm_plan :: Comm -> CInt -> CInt -> CInt
                         -> Int64 -> Int64 -> Int64 -> Int64
                         -> Ptr () -> Ptr () -> Ptr () -> CInt -> Ptr Plan   -- A B C root plan(out)
                         -> IO Status

foreign import ccall unsafe "accblas_mg_gemm_run"
  c_accblas_mg_gemm_run :: Plan -> Ptr () -> Ptr () -> IO Status          -- plan stream alpha(host)

foreign import ccall safe "accblas_mg_gemm_plan_destroy"
  c_accblas_mg_gemm_plan_destroy :: Plan -> IO Status

-- y = alpha * op(A) * x on row-sharded A; every rank receives the whole y
foreign import ccall unsafe "accblas_mg_gemv"
  c_accblas_mg_gemv :: Comm -> Ptr () -> CInt -> CInt
                    -> Int64 -> Int64 -> Ptr ()                           -- rows_local colsA alpha(host)
                    -> Ptr () -> Ptr () -> Int64 -> Ptr ()                -- A_local x |x| y_full
                    -> IO Status

foreign import ccall unsafe "accblas_mg_dot"
  c_accblas_mg_dot :: Comm -> Ptr () -> CInt -> CInt -> Int64             -- comm stream dtype conj_x n_local
                   -> Ptr () -> Ptr () -> Ptr () -> IO Status             -- x_local y_local out(device)

-- the library's private scratch pool (INTEGRATION.md section 4)
foreign import ccall unsafe "accblas_trim_workspace"
  c_accblas_trim_workspace :: IO Status

-- most recent CUDA failure of ANY thread, copied into the caller's buffer (FFI calls may migrate between OS threads)
foreign import ccall unsafe "accblas_last_cuda_error_r"
  c_accblas_last_cuda_error_r :: CString -> CSize -> IO CSize

-- per-thread kernel-choice knobs (accblas_tune_t); e.g. `c_accblas_set_tuning 7 0 >> c_accblas_set_tuning 9 0`
-- pins complex products to 4M.  Haskell FFI calls may hop between OS threads: prefer ACCBLAS_TUNE in the
-- environment (INTEGRATION.md) unless the caller runs in a bound thread.
foreign import ccall unsafe "accblas_set_tuning"
  c_accblas_set_tuning :: CInt -> CInt -> IO ()

foreign import ccall unsafe "accblas_status_string"
  c_accblas_status_string :: CInt -> IO CString

-- | The reference surfaces cuBLAS failures as exceptions from the binding; do the same.
checkStatus :: String -> Status -> IO ()
checkStatus loc st =
  when (st /= 0) $ do
    msg <- peekCString =<< c_accblas_status_string st
    throwIO (ErrorCall (loc ++ ": " ++ msg))
