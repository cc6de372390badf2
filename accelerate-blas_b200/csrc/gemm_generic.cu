// gemm_generic.cu -- shape-agnostic SIMT gemm: C = alpha * op(A) * op(B) for the four
// element types, any extents (odd, unaligned, tiny), N/T/H on either operand.
//
// This is the path for problems the tensor-core kernels do not take (row pitch not a
// multiple of 16 bytes, very small extents -- e.g. the 1..64 shapes of the reference's
// property tests, test/Test/BLAS/Level3.hs:68, or the k = 1 outer product `><`,
// LinearAlgebra.hs:73-78).  Replaces BLAS.?gemm at LLVM/PTX/Level3.hs:88-106 for those.
//
// Tiling: 64x64 output tile per 256-thread CTA, BK = 16, 4x4 register micro-tile,
// operands staged k-major in shared memory; global reads are coalesced along the
// stored row whichever op is requested.
#include "common.cuh"
#include <algorithm>

namespace accblas {
namespace {

constexpr int BM = 64, BN = 64, BK = 16, TM = 4, TN = 4, NT = 256;

// Read element (r, c) of a stored rows x cols row-major matrix with row pitch ld, 0 outside.
template <typename T>
__device__ __forceinline__ T ld_guard(const T *__restrict__ p, int64_t rows, int64_t cols, int64_t ld, int64_t r, int64_t c) {
    return (r < rows && c < cols) ? p[r * ld + c] : Scalar<T>::zero();
}

template <typename T, bool TA, bool TB>
__global__ void __launch_bounds__(NT)
gemm_generic_kernel(int64_t m, int64_t n, int64_t k, int64_t rowsA, int64_t colsA, int64_t rowsB,
                    int64_t colsB, int64_t lda, int64_t ldb, int64_t ldc, bool conjA, bool conjB, Alpha<T> alpha_,
                    const T *__restrict__ A, const T *__restrict__ B, T *C, BetaC<T> beta, unsigned tiles_n,
                    int64_t strideA, int64_t strideB, int64_t strideC)
{
    A += (int64_t)blockIdx.z * strideA;                        // strided batch: one product per grid plane
    B += (int64_t)blockIdx.z * strideB;
    C += (int64_t)blockIdx.z * strideC;
    const T alpha = alpha_.get();
    __shared__ T As[BK][BM + 1];
    __shared__ T Bs[BK][BN + 1];
    const int tid = threadIdx.x;
    const int tx = tid % 16, ty = tid / 16;
    // tiles are flattened onto gridDim.x (2^31 - 1 blocks): no 65535-row-tile limit on m
    const int64_t m0 = (int64_t)(blockIdx.x / tiles_n) * BM, n0 = (int64_t)(blockIdx.x % tiles_n) * BN;

    T acc[TM][TN];
#pragma unroll
    for (int i = 0; i < TM; ++i)
#pragma unroll
        for (int j = 0; j < TN; ++j) acc[i][j] = Scalar<T>::zero();

    for (int64_t k0 = 0; k0 < k; k0 += BK) {
        // ---- stage op(A)[m0.., k0..] as As[kk][mm] ----
#pragma unroll
        for (int e = tid; e < BM * BK; e += NT) {
            int mm, kk;
            T v;
            if (!TA) { kk = e % BK; mm = e / BK; v = ld_guard(A, rowsA, colsA, lda, m0 + mm, k0 + kk); }
            else     { mm = e % BM; kk = e / BM; v = ld_guard(A, rowsA, colsA, lda, k0 + kk, m0 + mm); }
            if (conjA) v = conj_s(v);
            As[kk][mm] = v;
        }
        // ---- stage op(B)[k0.., n0..] as Bs[kk][nn] ----
#pragma unroll
        for (int e = tid; e < BN * BK; e += NT) {
            int nn, kk;
            T v;
            if (!TB) { nn = e % BN; kk = e / BN; v = ld_guard(B, rowsB, colsB, ldb, k0 + kk, n0 + nn); }
            else     { kk = e % BK; nn = e / BK; v = ld_guard(B, rowsB, colsB, ldb, n0 + nn, k0 + kk); }
            if (conjB) v = conj_s(v);
            Bs[kk][nn] = v;
        }
        __syncthreads();
#pragma unroll
        for (int kk = 0; kk < BK; ++kk) {
            T a[TM], b[TN];
#pragma unroll
            for (int i = 0; i < TM; ++i) a[i] = As[kk][ty + 16 * i];
#pragma unroll
            for (int j = 0; j < TN; ++j) b[j] = Bs[kk][tx + 16 * j];
#pragma unroll
            for (int i = 0; i < TM; ++i)
#pragma unroll
                for (int j = 0; j < TN; ++j) acc[i][j] = fma_acc(a[i], b[j], acc[i][j]);
        }
        __syncthreads();
    }
#pragma unroll
    for (int i = 0; i < TM; ++i) {
        const int64_t r = m0 + ty + 16 * i;
        if (r >= m) continue;
#pragma unroll
        for (int j = 0; j < TN; ++j) {
            const int64_t c = n0 + tx + 16 * j;
            if (c < n) {
                T v = mul_s(alpha, acc[i][j]);                 // beta = 0 (the reference's only mode): C is never read
                if (beta.on()) v = beta.add(v, r, c);
                C[r * ldc + c] = v;
            }
        }
    }
}

template <typename T>
int launch_typed(const GemmArgs &g)
{
    const bool ta = g.opA != ACCBLAS_OP_N, tb = g.opB != ACCBLAS_OP_N;
    const bool ca = Scalar<T>::cplx && g.opA == ACCBLAS_OP_H;
    const bool cb = Scalar<T>::cplx && g.opB == ACCBLAS_OP_H;
    const Alpha<T> alpha = make_alpha<T>(g.alpha, g.alpha_dev);
    const int64_t batch = g.batch > 1 ? g.batch : 1;
    const int64_t tiles_n = (g.n + BN - 1) / BN, tiles = tiles_n * ((g.m + BM - 1) / BM);
    if (tiles > 0x7fffffff || batch > 65535) return ACCBLAS_ERR_SHAPE;   // callers chunk larger batches
    dim3 grid((unsigned)tiles, 1, (unsigned)batch);
    const BetaC<T> beta = make_beta<T>(g);
    const T *A = static_cast<const T *>(g.A), *B = static_cast<const T *>(g.B);
    T *C = static_cast<T *>(g.C);
#define GO(TA_, TB_)                                                                             \
    gemm_generic_kernel<T, TA_, TB_><<<grid, NT, 0, g.stream>>>(g.m, g.n, g.k, g.rowsA, g.colsA,   \
                                                                g.rowsB, g.colsB, g.lda, g.ldb, g.ldc, ca, cb,  \
                                                                alpha, A, B, C, beta, (unsigned)tiles_n,              \
                                                                g.strideA, g.strideB, g.strideC)
    if (!ta && !tb) GO(false, false);
    else if (!ta && tb) GO(false, true);
    else if (ta && !tb) GO(true, false);
    else GO(true, true);
#undef GO
    return check_launch("gemm_generic_kernel");
}

// ---------------------------------------------------------------------------------------
// k = 1: the outer product `x >< y = (len x, 1) <> (1, len y)` (LinearAlgebra.hs:73-78) and every
// other product with inner extent 1.  C[i][j] = (alpha * a_i) * b_j is a pure store stream:
// m * n * sizeof(T) bytes, no reuse worth a tile.  Each thread owns a 16-byte column vector and
// walks ROWS_PER_CTA rows; a_i, b_j come through the read-only path (b_j is re-used from registers
// across the rows).  Element strides cover all op / pitch cases: op(A)[i] = A[i * sa], op(B)[j] = B[j * sb].
// ---------------------------------------------------------------------------------------
constexpr int R1_THREADS = 256, R1_ROWS = 8;

template <typename T>
__global__ void __launch_bounds__(R1_THREADS)
gemm_rank1_kernel(int64_t m, int64_t n, int64_t sa, int64_t sb, int64_t ldc, bool conjA, bool conjB, Alpha<T> alpha_,
                  const T *__restrict__ A, const T *__restrict__ B, T *C, BetaC<T> beta, unsigned chunks_n, bool vec_ok)
{
    const T alpha = alpha_.get();
    constexpr int VEC = 16 / sizeof(T);
    // (row group, column chunk) flattened onto gridDim.x: no 65535 limit on m / R1_ROWS
    const int64_t j0 = ((int64_t)(blockIdx.x % chunks_n) * R1_THREADS + threadIdx.x) * VEC;
    const int64_t i0 = (int64_t)(blockIdx.x / chunks_n) * R1_ROWS;
    if (j0 >= n) return;
    T b[VEC];
#pragma unroll
    for (int e = 0; e < VEC; ++e) {
        b[e] = j0 + e < n ? __ldg(B + (j0 + e) * sb) : Scalar<T>::zero();
        if (conjB) b[e] = conj_s(b[e]);
    }
    const bool full = vec_ok && j0 + VEC <= n;
#pragma unroll
    for (int r = 0; r < R1_ROWS; ++r) {
        const int64_t i = i0 + r;
        if (i >= m) break;
        T a = __ldg(A + i * sa);
        if (conjA) a = conj_s(a);
        a = mul_s(alpha, a);                                   // (alpha * a) * b, BLAS/Level3.hs:89
        T *crow = C + i * ldc + j0;
        if (full) {
            union { int4 raw; T e[VEC]; } v;
#pragma unroll
            for (int e = 0; e < VEC; ++e) {
                v.e[e] = mul_s(a, b[e]);
                if (beta.on()) v.e[e] = beta.add(v.e[e], i, j0 + e);
            }
            *reinterpret_cast<int4 *>(crow) = v.raw;
        } else {
#pragma unroll
            for (int e = 0; e < VEC; ++e)
                if (j0 + e < n) {
                    T w = mul_s(a, b[e]);
                    if (beta.on()) w = beta.add(w, i, j0 + e);
                    crow[e] = w;
                }
        }
    }
}

template <typename T>
int launch_rank1(const GemmArgs &g)
{
    constexpr int VEC = 16 / sizeof(T);
    const Alpha<T> alpha = make_alpha<T>(g.alpha, g.alpha_dev);
    const bool ca = Scalar<T>::cplx && g.opA == ACCBLAS_OP_H, cb = Scalar<T>::cplx && g.opB == ACCBLAS_OP_H;
    // op(A) is m x 1: stored (m x 1, pitch lda) for N, stored (1 x m) for T/H
    const int64_t sa = g.opA == ACCBLAS_OP_N ? g.lda : 1;
    // op(B) is 1 x n: stored (1 x n) for N, stored (n x 1, pitch ldb) for T/H
    const int64_t sb = g.opB == ACCBLAS_OP_N ? 1 : g.ldb;
    const bool vec_ok = (reinterpret_cast<uintptr_t>(g.C) & 15u) == 0 && (g.ldc % VEC) == 0;
    const int64_t chunks_n = (g.n + (int64_t)R1_THREADS * VEC - 1) / ((int64_t)R1_THREADS * VEC);
    const int64_t blocks = chunks_n * ((g.m + R1_ROWS - 1) / R1_ROWS);
    if (blocks > 0x7fffffff) return gemm_generic(g);
    gemm_rank1_kernel<T><<<(unsigned)blocks, R1_THREADS, 0, g.stream>>>(g.m, g.n, sa, sb, g.ldc, ca, cb, alpha,
                                                            static_cast<const T *>(g.A), static_cast<const T *>(g.B),
                                                            static_cast<T *>(g.C), make_beta<T>(g), (unsigned)chunks_n, vec_ok);
    return check_launch("gemm_rank1_kernel");
}

}  // namespace

int gemm_rank1(const GemmArgs &g)
{
    switch (g.dtype) {
    case ACCBLAS_F32: return launch_rank1<float>(g);
    case ACCBLAS_F64: return launch_rank1<double>(g);
    case ACCBLAS_C32: return launch_rank1<float2>(g);
    default:          return launch_rank1<double2>(g);
    }
}

int gemm_generic(const GemmArgs &g)
{
    switch (g.dtype) {
    case ACCBLAS_F32: return launch_typed<float>(g);
    case ACCBLAS_F64: return launch_typed<double>(g);
    case ACCBLAS_C32: return launch_typed<float2>(g);
    default:          return launch_typed<double2>(g);
    }
}


// ---------------------------------------------------------------------------------------------------------------
// split-K (gemm_dmma.cu, gemm_mma_tf32.cu): the sum of the K slices' alpha-scaled partial products, in slice order
//     C[r][c] = sum_s P[s][r][c],   P = g.ksplit_ws (g.ksplit planes of m x n, row pitch n),  C with row pitch g.ldc
// ---------------------------------------------------------------------------------------------------------------
namespace {
template <typename T>
__global__ void __launch_bounds__(256) splitk_reduce_kernel(const T *__restrict__ P, int slices, int64_t m, int64_t n, T *__restrict__ C, int64_t ldc)
{
    const int64_t total = m * n;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
        T acc = P[i];
        for (int sl = 1; sl < slices; ++sl) acc = add_s(acc, P[(int64_t)sl * total + i]);
        const int64_t r = i / n;
        C[r * ldc + (i - r * n)] = acc;
    }
}
template <typename T> int launch_splitk_reduce(const GemmArgs &g)
{
    const int64_t total = g.m * g.n;
    const unsigned blocks = (unsigned)std::min<int64_t>((total + 255) / 256, 4 * (int64_t)sm_count());
    splitk_reduce_kernel<T><<<blocks, 256, 0, g.stream>>>(static_cast<const T *>(g.ksplit_ws), g.ksplit, g.m, g.n, static_cast<T *>(g.C), g.ldc);
    return check_launch("splitk_reduce_kernel");
}
}  // namespace

int splitk_reduce(const GemmArgs &g)
{
    switch (g.dtype) {
    case ACCBLAS_F32: return launch_splitk_reduce<float>(g);
    case ACCBLAS_F64: return launch_splitk_reduce<double>(g);
    case ACCBLAS_C32: return launch_splitk_reduce<float2>(g);
    default:          return launch_splitk_reduce<double2>(g);
    }
}

}  // namespace accblas
