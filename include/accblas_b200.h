/*
 * accblas_b200.h -- C ABI of libaccblas_b200.so
 *
 * Hand-written sm_100a (B200) kernels for the dense-product hot path of
 * tmcdonell/accelerate-blas: gemm, gemv, dotu/dotc over Float, Double,
 * Complex Float, Complex Double with the N / T / H operand flags.
 *
 * The library is the callee of the reference's llvm-ptx ForeignAcc hooks.  Every entry
 * point below names the reference interface it replaces (paths relative to the
 * reference tree, src/Data/Array/Accelerate/Numeric/LinearAlgebra/...).
 *
 * Contract:
 *   - stateless and re-entrant: no handle (the reference's withBLAS / cuBLAS handle
 *     cache, LLVM/PTX/Context.hs:45-75, disappears); safe on concurrent streams/devices;
 *     the CUDA device current on the calling thread is the one used;
 *   - stream ordered: ALL device work, including internal workspace allocation
 *     (cudaMallocAsync / cudaFreeAsync), is enqueued on the caller's stream, because the
 *     reference tracks buffer lifetimes with an event recorded on that stream right after
 *     the call (LLVM/PTX/Base.hs:94-98);
 *   - row-major, dense, tight leading dimensions (lda = colsA, ldb = colsB, ldc = n:
 *     LLVM/PTX/Level3.hs:64-65,91), stored shape is PRE-op, complex = interleaved (re,im)
 *     (Type.hs:35-36);
 *   - beta = 0 and the output is never read: the reference hands over an UNINITIALISED
 *     buffer (LLVM/PTX/Level3.hs:80,90; LLVM/PTX/Level2.hs:98,108);
 *   - `H` on a real dtype behaves exactly as `T` (BLAS/Level3.hs:97-100);
 *   - alpha is read from HOST memory at call time (the reference fetches it with
 *     indexRemote and passes a host pointer, LLVM/PTX/Level3.hs:81,89);
 *   - every function returns an accblas_status_t; nothing throws, nothing falls back to
 *     the CPU.  A zero extent launches nothing (k == 0 zero-fills the output).
 */
#ifndef ACCBLAS_B200_H
#define ACCBLAS_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define ACCBLAS_B200_VERSION 100 /* 0.1.0 */

/* NumericR -- Type.hs:32-36 */
typedef enum {
    ACCBLAS_F32 = 0, /* NumericRfloat32   : Float                      4 B/element  */
    ACCBLAS_F64 = 1, /* NumericRfloat64   : Double                     8 B/element  */
    ACCBLAS_C32 = 2, /* NumericRcomplex32 : Vec2 Float  (re,im)        8 B/element  */
    ACCBLAS_C64 = 3  /* NumericRcomplex64 : Vec2 Double (re,im)       16 B/element  */
} accblas_dtype_t;

/* Transpose -- Type.hs:78-90 (encodeTranspose, LLVM/PTX/Base.hs:44-47) */
typedef enum {
    ACCBLAS_OP_N = 0, /* as stored                        */
    ACCBLAS_OP_T = 1, /* implicitly transposed            */
    ACCBLAS_OP_H = 2  /* conjugated and transposed        */
} accblas_op_t;

typedef enum {
    ACCBLAS_OK = 0,
    ACCBLAS_ERR_DTYPE = 1,     /* dtype not in accblas_dtype_t                          */
    ACCBLAS_ERR_OP = 2,        /* op not in accblas_op_t                                */
    ACCBLAS_ERR_SHAPE = 3,     /* negative extent, inner extents differ, or |x| wrong   */
    ACCBLAS_ERR_NULL = 4,      /* null pointer with a non-empty extent                  */
    ACCBLAS_ERR_CUDA = 5,      /* a CUDA runtime / driver call failed (cudaGetLastError)*/
    ACCBLAS_ERR_ARCH = 6,      /* device is not compute capability 10.x                 */
    ACCBLAS_ERR_WORKSPACE = 7, /* caller-supplied workspace too small / misaligned      */
    ACCBLAS_ERR_ALIAS = 8      /* output overlaps an input                              */
} accblas_status_t;

/* stream: a cudaStream_t passed as an opaque pointer (NULL = the legacy default stream). */
typedef void *accblas_stream_t;

/* ------------------------------------------------------------------------------------
 * Level 3.   C (m x n) = alpha * op(A) * op(B)
 *
 * Replaces the four BLAS.{s,d,c,z}gemm calls of gemm'  LLVM/PTX/Level3.hs:88-106
 * (and, for the complex-H gemv route, as_gemm         LLVM/PTX/Level2.hs:57-69).
 *   (m,k) = opA == N ? (rowsA,colsA) : (colsA,rowsA)   LLVM/PTX/Level3.hs:57-59
 *   n     = opB == N ? colsB : rowsB                   LLVM/PTX/Level3.hs:60-62
 * Unlike the reference, the inner extents ARE compared (ACCBLAS_ERR_SHAPE).
 * alpha_host points at one element of `dtype` in host memory.
 * ---------------------------------------------------------------------------------- */
int accblas_gemm(accblas_stream_t stream, int dtype, int opA, int opB,
                 int64_t rowsA, int64_t colsA, int64_t rowsB, int64_t colsB,
                 const void *alpha_host, const void *A, const void *B, void *C);

/* The same product on WINDOWS of larger row-major arrays: lda, ldb, ldc are row pitches in
 * elements (>= the window's column count; 0 = tight).  The reference always passes tight
 * pitches (lda = colsA, ldb = colsB, ldc = n, LLVM/PTX/Level3.hs:64-65,91) because Accelerate
 * arrays are dense; this entry point is what accblas_gemm_host's upload/multiply/download
 * wavefront and the multi-GPU row panels are built from, and what a caller holding a slice
 * (Accelerate `slice`/`backpermute` fused away) can use without materialising it. */
int accblas_gemm_ld(accblas_stream_t stream, int dtype, int opA, int opB,
                    int64_t rowsA, int64_t colsA, int64_t rowsB, int64_t colsB,
                    const void *alpha_host, const void *A, int64_t lda, const void *B, int64_t ldb,
                    void *C, int64_t ldc);

/* The extended product  D = alpha * op(A) * op(B) + beta * Cin  on windows (pitches as in accblas_gemm_ld).
 * The reference never needs it -- its hook passes beta = 0 into an uninitialised output (LLVM/PTX/Level3.hs:80,90) --
 * but everything built ON TOP of the hot path does: the K-step loop of the 2-D (SUMMA) sharding, the K-chunked
 * products that overlap the broadcast of B in accblas_mg_gemm, and callers that fuse `zipWith (+) c (a <> b)`.
 *   beta_host : one element of `dtype` in host memory; NULL or a zero value => Cin is never read (may be NULL);
 *   Cin, ldcin: m x n, row pitch in elements (0 = ldc).  Cin may be C itself (same pointer AND pitch: in place) or
 *               an array that does not overlap C -- e.g. a local accumulator while C lives in a peer GPU's HBM.
 * Every kernel family applies beta in its own epilogue (no extra pass); k == 0 gives D = beta * Cin. */
int accblas_gemm_ex(accblas_stream_t stream, int dtype, int opA, int opB,
                    int64_t rowsA, int64_t colsA, int64_t rowsB, int64_t colsB,
                    const void *alpha_host, const void *A, int64_t lda, const void *B, int64_t ldb,
                    const void *beta_host, const void *Cin, int64_t ldcin, void *C, int64_t ldc);

/* `batch` independent products with identical shapes and flags: operand b of the batch lives at
 * base + b * stride (strides in ELEMENTS; strideA / strideB = 0 re-uses one operand for every product;
 * strideC >= m * n).  The reference has no batched operator (Accelerate would express it as a fold over a
 * rank-3 array and lose the foreign hook); this is SURVEY.md section 8(f) rank 3.
 *   - tiny products (< 48^3): one SIMT launch, the batch on gridDim.z;
 *   - products of the warp-level tensor families (f64 / c64 on DMMA, f32 / c32 on mma.sync: everything that does not fill
 *     the GPU by itself): ONE launch -- the batch is folded into the grid (blockIdx -> product, tile) and into a third
 *     dimension of the TMA tensor maps, so the tiles of all products share the waves and the tile choice sees
 *     tiles x batch.  Needs es * stride % 16 == 0 for the three strides (TMA plane stride, vector stores);
 *   - anything else (products large enough for the tcgen05 / 3M / persistent kernels, odd strides): enqueued back to
 *     back on `stream`. */
int accblas_gemm_strided_batched(accblas_stream_t stream, int dtype, int opA, int opB,
                                 int64_t rowsA, int64_t colsA, int64_t rowsB, int64_t colsB,
                                 const void *alpha_host, const void *A, int64_t strideA,
                                 const void *B, int64_t strideB, void *C, int64_t strideC, int64_t batch);

/* alpha in DEVICE memory, read by the kernel: the reference turns `alpha :: Exp e` into a device
 * Scalar (BLAS/Level3.hs:64) and then copies it back with a blocking indexRemote before every call
 * (LLVM/PTX/Level3.hs:81; LLVM/PTX/Level2.hs:99) because cuBLAS is used in host-pointer mode.  These
 * variants take that device Scalar as it is: no device-to-host copy, no stream synchronisation. */
int accblas_gemm_alpha_dev(accblas_stream_t stream, int dtype, int opA, int opB,
                           int64_t rowsA, int64_t colsA, int64_t rowsB, int64_t colsB,
                           const void *alpha_dev, const void *A, const void *B, void *C);
int accblas_gemv_alpha_dev(accblas_stream_t stream, int dtype, int opA, int64_t rowsA, int64_t colsA,
                           const void *alpha_dev, const void *A, const void *x, int64_t lenx, void *y);

/* Bytes of stream-ordered scratch the MAIN kernel family of accblas_gemm takes from the library's memory pool for
 * this problem: the hi / lo planes of the 3xTF32 path, the 3M planes, the stream-K accumulator slots (0 for the direct
 * kernels).  Informational, and a lower bound: it leaves out the two small extras a call may add -- the split-K
 * partial planes of a small product spread over idle SMs (S * m * n elements, S <= 16, products of at most a few
 * hundred rows and columns) and the re-pitched copy of an operand off the 16-byte grid (that operand's size with
 * its pitch rounded up).  accblas_workspace_cached_bytes() after a call shows what was really taken. */
size_t accblas_gemm_workspace_bytes(int dtype, int opA, int opB,
                                    int64_t rowsA, int64_t colsA, int64_t rowsB, int64_t colsB);

/* ------------------------------------------------------------------------------------
 * Level 2.   y = alpha * op(A) * x,   |y| = opA == N ? rowsA : colsA
 *
 * Replaces the four BLAS.{s,d,c,z}gemv calls of as_gemv  LLVM/PTX/Level2.hs:105-124 and
 * the gemm detour for complex H (gemv' dispatch, LLVM/PTX/Level2.hs:48-54): all three
 * ops of all four dtypes are served by native row-major gemv kernels.
 * lenx must equal (opA == N ? colsA : rowsA).
 * ---------------------------------------------------------------------------------- */
int accblas_gemv(accblas_stream_t stream, int dtype, int opA,
                 int64_t rowsA, int64_t colsA,
                 const void *alpha_host, const void *A, const void *x, int64_t lenx, void *y);

/* ------------------------------------------------------------------------------------
 * Level 1.   out = sum_i x_i * y_i            (conj_x == 0: dotu, BLAS/Level1.hs:73-74)
 *            out = sum_i conj(x_i) * y_i      (conj_x != 0: dotc, BLAS/Level1.hs:85-92)
 *
 * NEW hook: the reference has no foreign dot (it is `fold (+) 0 (zipWith (*) xs ys)`);
 * INTEGRATION.md shows the foreignAcc alternative to add at BLAS/Level1.hs:74,91-92.
 * n = min(|x|,|y|) (zipWith intersects extents).  The result (one element of `dtype`)
 * is written to DEVICE memory at out_dev, like the reference's Acc (Scalar e), with a
 * fixed summation order (bit-reproducible run to run).  n == 0 writes 0.
 * conj_x is ignored for real dtypes.
 * ---------------------------------------------------------------------------------- */
int accblas_dot(accblas_stream_t stream, int dtype, int conj_x, int64_t n,
                const void *x, const void *y, void *out_dev);
/* The same product with COMPENSATED arithmetic (the reference's Numeric/Sum.hs): every product is
 * split exactly into (p, e) with one FMA, running sums are TwoSums whose rounding errors are kept beside them, and
 * partial sums of threads, warps and CTAs are merged with the reference's Kahan-Babuska-Neumaier combine (`kbnAdd`,
 * Numeric/Sum.hs:119-125); out = s + c rounded once.  As accurate as a dot evaluated in twice the working precision:
 * |out - exact| <= eps |exact| + O(n^2 eps^2) sum_i |x_i y_i| with eps of DOUBLE for every dtype (float operands are
 * multiplied exactly and summed in double).  Same HBM traffic as accblas_dot, same fixed summation tree. */
int accblas_dot_compensated(accblas_stream_t stream, int dtype, int conj_x, int64_t n,
                            const void *x, const void *y, void *out_dev);
/* accblas_dot for callers that cannot let the library allocate device memory.  No counterpart in the reference
 * (which has no foreign dot at all); it exists for hosts like these:
 *   - a runtime that budgets the whole of HBM through its own memory manager (Accelerate's llvm-ptx backend keeps a
 *     table of every device array and evicts by it): memory a foreign function takes behind its back is memory the
 *     manager still believes free, and the next allocation of either side can fail;
 *   - a process whose device memory has all been handed to an arena or caching allocator, where cudaMalloc simply
 *     returns out-of-memory;
 *   - latency-sensitive callers with several streams in flight: cudaMalloc / cudaFree synchronise the device.
 * accblas_dot keeps 33 KiB from cudaMalloc per (host thread, device, stream) or takes its partial sums from the
 * library's memory pool.  This variant does neither: the per-CTA partial sums go to a fixed array inside the library
 * image (ACCBLAS_DOT_STATIC_SLOTS slots of 32 bytes = 32 KiB; every GPU that loads the library holds it until the
 * process exits).  It calls no cudaMalloc, cudaFree, pool or other allocating routine; x, y and out_dev are still the
 * caller's.  Same arguments, same result BITS as accblas_dot with default tuning (same grid, same summation tree);
 * two launches instead of one.
 *
 * Limits (ACCBLAS_ERR_WORKSPACE, with a sentence in accblas_last_cuda_error): n <= ACCBLAS_DOT_STATIC_MAX_N, and the
 * launch must need at most ACCBLAS_DOT_STATIC_SLOTS CTAs (it uses two per SM: 296 on a B200).  Use accblas_dot beyond.
 *
 * First call on each device: looks the array's device address up (cudaGetSymbolAddress, which may load the module on
 * that device) with that device current -- make that first call outside any stream capture.  Later calls do no set-up.
 *
 * Every call on a device shares the one array, whatever the stream or host thread.  Calls on the same stream are
 * ordered and safe.  Calls that can run on the GPU at the same time (different streams, no event between them)
 * overwrite each other's partial sums and return garbage: serialise them, or use accblas_dot for all but one stream.
 * Nothing in the array has to be zero between calls: each call writes every slot it later reads. */
#define ACCBLAS_DOT_STATIC_SLOTS 1024
#define ACCBLAS_DOT_STATIC_MAX_N ((int64_t)1 << 32)
int accblas_dot_static(accblas_stream_t stream, int dtype, int conj_x, int64_t n,
                       const void *x, const void *y, void *out_dev);

/* ------------------------------------------------------------------------------------
 * The other Level-1 reductions of the reference (BLAS/Level1.hs).  Like dot they are plain
 * Accelerate folds there -- NEW hooks, one pass over HBM each, result in DEVICE memory.
 *
 *   sdot z xs ys  (Level1.hs:44-62)   out = z + sum_i x_i * y_i with every product and the sum in
 *                 DOUBLE precision whatever the element type, rounded once to `dtype` at the end
 *                 (`dsdot` / `zsdot`).  z_host: one element of `dtype` in host memory.
 *   asum xs       (Level1.hs:100-109) out = sum_i |x_i|, complex: |re| + |im|.  The result has the
 *                 element's BASE type (NumericBaseT: float for F32/C32, double for F64/C64).
 *   amax / amin   (Level1.hs:115-143) 0-based index (int64, the reference's `Int`) of the element
 *                 of largest / smallest magnitude (same magnitude as asum, computed in the
 *                 element's precision).  On ties the LAST such index is returned: the reference's
 *                 `cmp ix iy = snd ix > snd iy ? ix : iy` keeps the right operand (cblas_i?amax
 *                 returns the first).  n == 0 is ACCBLAS_ERR_SHAPE (fold1 of an empty array).
 * ---------------------------------------------------------------------------------- */
int accblas_sdot(accblas_stream_t stream, int dtype, int64_t n, const void *z_host,
                 const void *x, const void *y, void *out_dev);
int accblas_asum(accblas_stream_t stream, int dtype, int64_t n, const void *x, void *out_dev);
int accblas_amax(accblas_stream_t stream, int dtype, int64_t n, const void *x, int64_t *out_dev);
int accblas_amin(accblas_stream_t stream, int dtype, int64_t n, const void *x, int64_t *out_dev);

/* ------------------------------------------------------------------------------------
 * Host-buffer entry points: what `run (A <> B)` does end to end in the reference --
 * `use` uploads the host arrays, the hook runs, the result is copied back
 * (test/Test/BLAS/Level3.hs:52-54; bench/Bench/Accelerate.hs:111-114).  A, B, x, y, C,
 * out are HOST pointers (pinned memory gives full PCIe rate, pageable works);
 * gemm overlaps upload, multiply and download as a wavefront over a block grid of C: step t
 * uploads row panel t of op(A) and column panel t of op(B), which releases the L-shaped strip
 * C[t, 0..t-1] + C[0..t, t] for multiplication -- or, when the product is compute-bound against the PCIe link, as K-panel
 * rank updates of the whole C followed by a row-panel finish with the downloads underneath (csrc/host_pipeline.cu).
 * gemv op N uploads A in row chunks and multiplies chunk i under the upload of chunk i + 1; gemv op T / H and the Level-1
 * routines are upload -> kernel -> download (their kernels are < 1 % of the transfer).  Synchronous: returns
 * when the result is in host memory.  `stream` orders the work like above.
 * ---------------------------------------------------------------------------------- */
int accblas_gemm_host(accblas_stream_t stream, int dtype, int opA, int opB,
                      int64_t rowsA, int64_t colsA, int64_t rowsB, int64_t colsB,
                      const void *alpha_host, const void *A_host, const void *B_host, void *C_host);
int accblas_gemv_host(accblas_stream_t stream, int dtype, int opA, int64_t rowsA, int64_t colsA,
                      const void *alpha_host, const void *A_host, const void *x_host, int64_t lenx,
                      void *y_host);
int accblas_dot_host(accblas_stream_t stream, int dtype, int conj_x, int64_t n,
                     const void *x_host, const void *y_host, void *out_host);

int accblas_sdot_host(accblas_stream_t stream, int dtype, int64_t n, const void *z_host,
                      const void *x_host, const void *y_host, void *out_host);
int accblas_asum_host(accblas_stream_t stream, int dtype, int64_t n, const void *x_host, void *out_host);
int accblas_amax_host(accblas_stream_t stream, int dtype, int64_t n, const void *x_host, int64_t *out_host);
int accblas_amin_host(accblas_stream_t stream, int dtype, int64_t n, const void *x_host, int64_t *out_host);

/* ------------------------------------------------------------------------------------
 * Peer memory (one process per GPU, NVLink / NVSwitch).  NEW: the reference has no multi-GPU
 * path; its only provision is one cuBLAS handle per CUDA context (LLVM/PTX/Context.hs:50-65).
 * BASELINE.json configs[4] shards gemm by rows of op(A)/C with the operands on a root GPU.
 * These entry points let a peer rank map the root's arrays, so that
 *   - the row panel of A is pulled by the copy engines (no SMs on either side), and
 *   - accblas_gemm is handed a C pointer INSIDE THE ROOT'S HBM: the epilogue's stores are
 *     the gather (fused, tile by tile, no collective kernel).
 * The handle is plain bytes: ship it with any transport (torch.distributed, MPI, a pipe).
 * ---------------------------------------------------------------------------------- */
typedef struct {
    unsigned char ipc[64]; /* cudaIpcMemHandle_t of the allocation                      */
    uint64_t offset;       /* byte offset of the exported pointer inside the allocation */
    uint64_t size;         /* bytes of the allocation                                   */
    int32_t device;        /* exporting process's device ordinal (informational)        */
    int32_t reserved;
} accblas_peer_handle_t;

/* Describe the cudaMalloc allocation that contains dev_ptr (call in the owning process). */
int accblas_peer_export(const void *dev_ptr, accblas_peer_handle_t *out);
/* Map it in another process on the same box; *mapped corresponds to the exporter's dev_ptr.
 * Mappings are cached per process and stay valid until accblas_peer_close_all(). */
int accblas_peer_open(const accblas_peer_handle_t *h, void **mapped);
int accblas_peer_close_all(void);
/* Pitched device-to-device copy on `stream` (copy engines; either side may be peer-mapped). */
int accblas_peer_copy2d(accblas_stream_t stream, void *dst, size_t dst_pitch, const void *src,
                        size_t src_pitch, size_t width_bytes, size_t rows);

/* ------------------------------------------------------------------------------------
 * Multi-GPU (one process per GPU of one NVLink / NVSwitch box).  NEW: the reference is single-GPU (one cuBLAS
 * handle per CUDA context, LLVM/PTX/Context.hs:50-65); BASELINE.json configs[4] and SURVEY.md section 8(e) define the
 * decomposition.  A host runtime with one OS process per device (e.g. one Accelerate `PTX` context each) calls these
 * the way it calls accblas_gemm; `comm` is created once per process.
 *
 *   accblas_mg_unique_id   rank 0: 128 bootstrap bytes; ship them to the other ranks with any transport.
 *   accblas_mg_init        collective; uses the CUDA device current on the calling thread.  Loads NCCL with dlopen
 *                          (libnccl.so.2 or $ACCBLAS_NCCL_LIB) -- the library itself links only cudart.
 *   accblas_mg_gemm        C = alpha * op(A) * op(B), operands and result on `root` (pass NULL elsewhere); rows of
 *                          op(A) / C are split over the ranks (accblas_mg_row_panel).  Peers PULL their A panel and B
 *                          from the root's HBM with the copy engines in K-chunks (accblas_mg_k_chunks), multiply
 *                          each chunk as it lands (beta = 1 accumulation, accblas_gemm_ex) and store the last chunk's
 *                          result straight into the root's C over NVLink (gather fused into the epilogue).  The
 *                          root's A, B, C must come from cudaMalloc (CUDA IPC).  Collective.  This one-shot form
 *                          plans, runs, WAITS for the stream and frees the peers' replicas on every call (it is
 *                          synchronous); loops should use the plan / run split below, whose run is stream-ordered:
 *                          when its work has completed on the root's stream, C is complete.
 *                          A failure on any rank while planning (an operand that is not a cudaMalloc allocation, a
 *                          mapping or replica that cannot be made) is reported on EVERY rank and no rank keeps a
 *                          plan; a failure of one rank's local work in run still passes the closing token, so the
 *                          other ranks return instead of waiting.
 *   accblas_mg_gemm_plan / _run / _plan_destroy
 *                          the same, split: the plan exchanges the IPC handles and allocates the peers' replicas
 *                          once, so that a loop over the same arrays pays no handle exchange or host
 *                          synchronisation per product.
 *   accblas_mg_gemv        y = alpha * op(A) * x with the ROWS of the stored A pre-sharded (A_local = rows_local x colsA).
 *                          op N: x (colsA) replicated, every rank receives the full y = the ranks' slices in rank
 *                          order.  The panels may differ in height and may be empty (accblas_mg_row_panel cuts whole
 *                          128-row tiles: m = 1000 on 8 ranks is 7 x 128 + 104, m = 256 leaves ranks 2..7 empty):
 *                          each rank's length travels with its deposit; y_full holds the SUM of all rows_local, and a
 *                          rank with no rows still makes the call.  op T / H: x = this rank's slice (rows_local);
 *                          every rank receives the full y (colsA), partials summed in rank order.
 *   accblas_mg_dot         dotu / dotc over pre-sharded slices; every rank receives the scalar in out_dev.
 *                          gemv / dot exchange through IPC-mapped buffers with flags (no collective kernel);
 *                          at most 1 MiB of y per rank (ACCBLAS_ERR_WORKSPACE beyond).  Calls on one communicator
 *                          must be issued in the same order by every rank, on streams ordered one after another.
 * ---------------------------------------------------------------------------------- */
#define ACCBLAS_MG_ID_BYTES 128
typedef struct accblas_mg_comm_s *accblas_mg_comm_t;
typedef struct accblas_mg_plan_s *accblas_mg_plan_t;
int accblas_mg_unique_id(void *id_out);
int accblas_mg_init(accblas_mg_comm_t *comm, int rank, int world, const void *id_bytes);
int accblas_mg_finalize(accblas_mg_comm_t comm);
int accblas_mg_rank(accblas_mg_comm_t comm);
int accblas_mg_world(accblas_mg_comm_t comm);
/* this rank's panel of the m rows: 128-aligned contiguous panels, ranks past the end get rows = 0 */
int accblas_mg_row_panel(int64_t m, int world, int rank, int64_t *start, int64_t *rows);
/* the K-chunks of the pipelined product (k0[i], dk[i]); returns their number (<= cap) */
int accblas_mg_k_chunks(int64_t k, int64_t *k0_out, int64_t *dk_out, int cap);
int accblas_mg_gemm(accblas_mg_comm_t comm, accblas_stream_t stream, int dtype, int opA, int opB,
                    int64_t rowsA, int64_t colsA, int64_t rowsB, int64_t colsB,
                    const void *alpha_host, const void *A, const void *B, void *C, int root);
int accblas_mg_gemm_plan(accblas_mg_comm_t comm, int dtype, int opA, int opB,
                         int64_t rowsA, int64_t colsA, int64_t rowsB, int64_t colsB,
                         const void *A, const void *B, void *C, int root, accblas_mg_plan_t *plan);
int accblas_mg_gemm_run(accblas_mg_plan_t plan, accblas_stream_t stream, const void *alpha_host);
int accblas_mg_gemm_plan_destroy(accblas_mg_plan_t plan);
int accblas_mg_gemv(accblas_mg_comm_t comm, accblas_stream_t stream, int dtype, int opA,
                    int64_t rows_local, int64_t colsA, const void *alpha_host,
                    const void *A_local, const void *x, int64_t lenx, void *y_full);
int accblas_mg_dot(accblas_mg_comm_t comm, accblas_stream_t stream, int dtype, int conj_x, int64_t n_local,
                   const void *x_local, const void *y_local, void *out_dev);

/* ------------------------------------------------------------------------------------
 * Introspection / tuning.  None of these change results beyond the stated tolerance.
 * ---------------------------------------------------------------------------------- */
int accblas_version(void);                     /* ACCBLAS_B200_VERSION */
const char *accblas_status_string(int status); /* static string        */
/* cudaGetLastError text captured by the last ACCBLAS_ERR_CUDA on this thread ("" if none) */
const char *accblas_last_cuda_error(void);
/* The same for callers whose FFI calls hop between OS threads (GHC `unsafe` foreign imports): copies the text of
 * the most recent ACCBLAS_ERR_CUDA of ANY thread of the process into buf (NUL-terminated, truncated to cap);
 * returns the full length. */
size_t accblas_last_cuda_error_r(char *buf, size_t cap);
/* Name of the kernel family the dispatcher picks for this gemm problem, e.g.
 * "dgemm_dmma_tma", "sgemm_tf32x3_tcgen05", "sgemm_tf32x3_mma", "gemm_rank1_outer",
 * "gemm_simt_generic".  Static string. */
const char *accblas_gemm_kernel_name(int dtype, int opA, int opB,
                                     int64_t rowsA, int64_t colsA, int64_t rowsB, int64_t colsB);
/* Number of kernel launches this thread has enqueued through the library (bench.py's
 * gpu_launches). */
uint64_t accblas_launch_count(void);
/* ... and by every thread of the process (thread-migration-proof variant). */
uint64_t accblas_launch_count_total(void);

/* Workspace.  Stream-ordered scratch (3xTF32 planes, 3M planes, stream-K partials, re-pitched operands, reduction
 * partials, the host-buffer entry points' device copies) comes from a memory pool the library OWNS, one per device,
 * created on first use: the device's default pool -- which an Accelerate process shares with everything else
 * that calls cudaMallocAsync -- is never touched.  Freed blocks stay cached in the pool (a physical allocation
 * per call would cost ~0.2 ms); accblas_trim_workspace returns them to the driver, e.g. from the host
 * runtime's out-of-memory handler (Accelerate's allocator reclaims on OOM).  Current device only. */
int accblas_trim_workspace(void);
size_t accblas_workspace_cached_bytes(void);     /* bytes the current device's pool holds right now */
/* Force a kernel family for gemm on this thread: 0 = auto, 1 = generic SIMT,
 * 2 = tensor path only (error ACCBLAS_ERR_SHAPE if the problem does not qualify). */
void accblas_set_gemm_path(int path);

/* Per-thread tuning knobs for benchmarking the kernel variants against each other
 * (tools/perf.py, baseline/sidebyside).  value < 0 restores the automatic choice.  Results
 * stay within the stated tolerance whatever the setting. */
typedef enum {
    ACCBLAS_TUNE_DMMA_TILE = 0, /* f64 gemm CTA tile: 0 = 128x128, 1 = 64x128, 2 = 64x64, 3 = 32x32 */
    ACCBLAS_TUNE_ZMMA_TILE = 1, /* c64 gemm CTA tile: 0 = 128x64, 1 = 64x64, 2 = 32x32            */
    ACCBLAS_TUNE_SGEMM_FAMILY = 2, /* f32/c32 gemm: 1 = mma.sync path, 2 = tcgen05 path (if eligible)    */
    ACCBLAS_TUNE_SMMA_TILE = 3, /* mma.sync path CTA tile: 0 = 128x64 (c32: 64x64), 1 = 64x64, 2 = 32x32 */
    ACCBLAS_TUNE_HOST_PIPELINE = 4, /* accblas_gemm_host: 1 = wavefront even for small transfers, 0 = never */
    ACCBLAS_TUNE_RASTER_GROUP = 5,  /* f64 gemm: tile rows per rasterisation group (default 16)              */
    ACCBLAS_TUNE_HOST_PANELS = 6,   /* accblas_gemm_host: maximum panels per dimension of the wavefront     */
    ACCBLAS_TUNE_ZGEMM_3M = 7,      /* c64 gemm: 0 = always 4M in registers, 1 = 3M (three real products) whenever possible */
    ACCBLAS_TUNE_TC_TILE = 8,       /* tcgen05 f32/c32 gemm tile: 0 = 128x256, 1 = 128x128, 2 = 128x64, 3 = 256x256 on a CTA pair (cta_group::2) */
    ACCBLAS_TUNE_CGEMM_3M = 9,      /* c32 gemm on tcgen05: 0 = always the 4M real embedding, 1 = 3M whenever possible     */
    ACCBLAS_TUNE_SNAKE_K = 10,      /* f64 gemm: 0 = every tile walks K forwards (default: alternate per wave of CTAs)       */
    ACCBLAS_TUNE_STREAMK = 11,      /* f64 gemm: 0 = never the persistent kernel, 1 = whole-tile rounds + stream-K tail whenever
                                       possible, 2 = pure stream-K (no whole-tile rounds)                                          */
    ACCBLAS_TUNE_RELEASE = 12,      /* f64/c64 gemm stage release: 0 = fence.proxy.async + arrive everywhere, 2 = release delayed
                                       by one K-block (no fence needed) everywhere; default: delayed on the data-parallel 128x128
                                       f64 tile, fenced elsewhere (csrc/gemm_dmma.cu: release_mode)                                 */
    ACCBLAS_TUNE_MN_ONEBOX = 13,    /* f64 gemm: 0 = fetch an MN-major operand panel as 16 x 16 TMA boxes (round 1), default: as ONE
                                       3-D box per stage when the extent is a multiple of 16                                       */
    ACCBLAS_TUNE_MG_SELF_PEER = 14, /* accblas_mg_gemm_plan: 1 = the root also runs the peers' pull / K-chunk pipeline on its own
                                       arrays (test hook: exercises that path on a one-GPU box)                                    */
    ACCBLAS_TUNE_DOT_SHAPE = 15,    /* dot: CTA shape of the streaming pass, threads x 16-byte loads per operand per thread x CTAs
                                       per SM: 0 = 512x4x2 (default), 1 = 256x4x4, 2 = 256x8x2, 4 = 512x8x1,
                                       5 = 1024x4x1 (profiles/r02_dot_shapes.jsonl)                                                */
    ACCBLAS_TUNE_MG_DOT_FUSED = 16, /* accblas_mg_dot: 0 = local dot (two launches) + one exchange kernel, default: ONE launch whose
                                       last CTA sums the partials and runs the scalar all-reduce over the peers' buffers           */
    ACCBLAS_TUNE_SGEMM_DIRECT = 17, /* f32 gemm, operand base / pitch off TMA's 16-byte grid, k <= 768: 0 = copy into aligned scratch first
                                       (round 1), default: the mma.sync kernel loads its panels itself (one launch, no scratch)     */
    ACCBLAS_TUNE_DGEMM_DIRECT = 18, /* f64 gemm, operand off TMA's 16-byte grid (odd column count), at most one 32 x 32 tile per SM, k <= 384:
                                       0 = copy into aligned scratch first (round 1), default: the DMMA kernel loads its panels itself   */
    ACCBLAS_TUNE_HOST_SCHEDULE = 19, /* accblas_gemm_host: 0 = always the block wavefront, 1 = always K-panel rank updates + row-panel finish,
                                        default: K-panels when the product is compute-bound against the PCIe link (csrc/host_pipeline.cu) */
    ACCBLAS_TUNE_SPLITK = 20,       /* f64 / c64 gemm: 0 = never split K over idle SMs (round 1), default: products of at most half a wave of
                                       32 x 32 tiles with k >= 512 (c64: 256) run up to 16 K slices + a reduce kernel                 */
    ACCBLAS_TUNE_DOT_FUSED = 21,    /* dot: 0 = partial kernel + final kernel (round 1), default: one launch whose last CTA sums the partials
                                       (needs 33 KiB of device memory kept per thread and stream)                                  */
    ACCBLAS_TUNE_COUNT = 22
} accblas_tune_t;
/* Process-wide defaults for threads that never call this: environment variable
 * ACCBLAS_TUNE="knob=value,knob=value" (read once). */
void accblas_set_tuning(int knob, int value);

#ifdef __cplusplus
}
#endif
#endif /* ACCBLAS_B200_H */
