// dot.cu -- dotu / dotc as a grid-level reduction (HBM-bound: 2 * n * sizeof(elem) bytes).
//
// NEW foreign hook for BLAS/Level1.hs:73-74 (dotu) and :85-92 (dotc, conjugate on the
// FIRST vector); the reference evaluates `fold (+) 0 (zipWith (*) xs ys)` with the
// backend's generic fold.  Result stays on the device like the reference's Acc (Scalar e).
//
// Pass 1: a fixed grid of 2 CTAs per SM, 512 threads each, walks the vectors in chunks
//         of 512 x UNROLL 16-byte vectors; every thread keeps UNROLL x 2 128-bit streaming
//         loads in flight (L1::no_allocate), accumulates in registers, then
//         shuffle + shared-memory block reduce -> one double-precision partial per CTA.
// Pass 2: one CTA sums the partials in index order and rounds once to the element type.
// The summation tree depends only on (n, SM count): results are bit-reproducible.
#include "common.cuh"

namespace accblas {
namespace {

constexpr int DOT_THREADS = 512;
constexpr int DOT_UNROLL = 4;

// accumulation type: real parts are accumulated in double for the cross-thread stages
template <typename T> struct Acc;
template <> struct Acc<float>   { using type = double;  };
template <> struct Acc<double>  { using type = double;  };
template <> struct Acc<float2>  { using type = double2; };
template <> struct Acc<double2> { using type = double2; };

__device__ __forceinline__ double  widen(float v) { return (double)v; }
__device__ __forceinline__ double  widen(double v) { return v; }
__device__ __forceinline__ double2 widen(float2 v) { return make_double2(v.x, v.y); }
__device__ __forceinline__ double2 widen(double2 v) { return v; }
template <typename T> __device__ __forceinline__ T narrow(typename Acc<T>::type v);
template <> __device__ __forceinline__ float   narrow<float>(double v) { return (float)v; }
template <> __device__ __forceinline__ double  narrow<double>(double v) { return v; }
template <> __device__ __forceinline__ float2  narrow<float2>(double2 v) { return make_float2((float)v.x, (float)v.y); }
template <> __device__ __forceinline__ double2 narrow<double2>(double2 v) { return v; }

template <typename T> union Vec16 {
    int4 raw;
    T e[16 / sizeof(T)];
};

template <typename A>
__device__ __forceinline__ A block_sum(A v, A *smem)
{
    v = warp_sum(v);
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    if (lane == 0) smem[warp] = v;
    __syncthreads();
    if (warp == 0) {
        A z;
        if constexpr (sizeof(A) == 16) z = make_double2(0.0, 0.0); else z = 0.0;
        v = lane < (int)(blockDim.x >> 5) ? smem[lane] : z;
        v = warp_sum(v);
    }
    return v;   // valid in warp 0
}

template <typename T, bool CONJ, bool ALIGNED>
__global__ void __launch_bounds__(DOT_THREADS, 2)
dot_partial_kernel(int64_t n, const T *__restrict__ x, const T *__restrict__ y,
                   typename Acc<T>::type *__restrict__ partial)
{
    using A = typename Acc<T>::type;
    constexpr int VEC = 16 / sizeof(T);
    __shared__ A red[DOT_THREADS / 32];
    T acc[DOT_UNROLL];
#pragma unroll
    for (int u = 0; u < DOT_UNROLL; ++u) acc[u] = Scalar<T>::zero();

    if (ALIGNED) {
        const int64_t nv = n / VEC;
        constexpr int64_t CHUNK = (int64_t)DOT_THREADS * DOT_UNROLL;
        const int64_t nchunks = nv / CHUNK;
        const int4 *xv = reinterpret_cast<const int4 *>(x);
        const int4 *yv = reinterpret_cast<const int4 *>(y);
        for (int64_t c = blockIdx.x; c < nchunks; c += gridDim.x) {
            const int64_t base = c * CHUNK + threadIdx.x;
            Vec16<T> a[DOT_UNROLL], b[DOT_UNROLL];
#pragma unroll
            for (int u = 0; u < DOT_UNROLL; ++u) a[u].raw = ldg_stream(xv + base + u * DOT_THREADS);
#pragma unroll
            for (int u = 0; u < DOT_UNROLL; ++u) b[u].raw = ldg_stream(yv + base + u * DOT_THREADS);
#pragma unroll
            for (int u = 0; u < DOT_UNROLL; ++u)
#pragma unroll
                for (int e = 0; e < VEC; ++e)
                    acc[u] = fma_acc(CONJ ? conj_s(a[u].e[e]) : a[u].e[e], b[u].e[e], acc[u]);
        }
        // remainder: whole vectors past the last full chunk, then the scalar tail
        for (int64_t v = nchunks * CHUNK + (int64_t)blockIdx.x * DOT_THREADS + threadIdx.x; v < nv;
             v += (int64_t)gridDim.x * DOT_THREADS) {
            Vec16<T> a, b;
            a.raw = ldg_stream(xv + v);
            b.raw = ldg_stream(yv + v);
#pragma unroll
            for (int e = 0; e < VEC; ++e) acc[0] = fma_acc(CONJ ? conj_s(a.e[e]) : a.e[e], b.e[e], acc[0]);
        }
        if (blockIdx.x == 0 && threadIdx.x < (int)(n - nv * VEC)) {
            const int64_t i = nv * VEC + threadIdx.x;
            acc[1] = fma_acc(CONJ ? conj_s(x[i]) : x[i], y[i], acc[1]);
        }
    } else {
        for (int64_t i = (int64_t)blockIdx.x * DOT_THREADS + threadIdx.x; i < n;
             i += (int64_t)gridDim.x * DOT_THREADS)
            acc[0] = fma_acc(CONJ ? conj_s(x[i]) : x[i], y[i], acc[0]);
    }
    A s = widen(acc[0]);
#pragma unroll
    for (int u = 1; u < DOT_UNROLL; ++u) s = add_s(s, widen(acc[u]));
    s = block_sum(s, red);
    if (threadIdx.x == 0) partial[blockIdx.x] = s;
}

template <typename T>
__global__ void __launch_bounds__(DOT_THREADS)
dot_final_kernel(int nparts, const typename Acc<T>::type *__restrict__ partial, T *__restrict__ out)
{
    using A = typename Acc<T>::type;
    __shared__ A red[DOT_THREADS / 32];
    A s;
    if constexpr (sizeof(A) == 16) s = make_double2(0.0, 0.0); else s = 0.0;
    for (int i = threadIdx.x; i < nparts; i += DOT_THREADS) s = add_s(s, partial[i]);
    s = block_sum(s, red);
    if (threadIdx.x == 0) *out = narrow<T>(s);
}

template <typename T>
int dot_typed(cudaStream_t st, bool conj, int64_t n, const T *x, const T *y, T *out)
{
    using A = typename Acc<T>::type;
    if (n == 0) {   // fold (+) 0 over nothing
        ACC_CUDA(cudaMemsetAsync(out, 0, sizeof(T), st));
        return ACCBLAS_OK;
    }
    constexpr int VEC = 16 / sizeof(T);
    const bool aligned = ((reinterpret_cast<uintptr_t>(x) | reinterpret_cast<uintptr_t>(y)) & 15u) == 0;
    const int64_t work = aligned ? (n / VEC + DOT_THREADS - 1) / DOT_THREADS : (n + DOT_THREADS - 1) / DOT_THREADS;
    int grid = 2 * sm_count();
    if (work < grid) grid = (int)(work > 0 ? work : 1);
    A *partial = nullptr;
    ACC_CUDA(scratch_alloc(reinterpret_cast<void **>(&partial), sizeof(A) * (size_t)grid, st));
    const bool cj = conj && Scalar<T>::cplx;
#define GO(CJ, AL) dot_partial_kernel<T, CJ, AL><<<grid, DOT_THREADS, 0, st>>>(n, x, y, partial)
    if (aligned) { if (cj) GO(true, true); else GO(false, true); }
    else         { if (cj) GO(true, false); else GO(false, false); }
#undef GO
    int s = check_launch("dot_partial_kernel");
    if (s == ACCBLAS_OK) {
        dot_final_kernel<T><<<1, DOT_THREADS, 0, st>>>(grid, partial, out);
        s = check_launch("dot_final_kernel");
    }
    cudaError_t e = cudaFreeAsync(partial, st);
    if (s == ACCBLAS_OK && e != cudaSuccess) return cuda_fail(e, "cudaFreeAsync");
    return s;
}

}  // namespace

int dot_dispatch(cudaStream_t s, int dtype, int conj_x, int64_t n, const void *x, const void *y, void *out)
{
    switch (dtype) {
    case ACCBLAS_F32: return dot_typed<float>(s, false, n, (const float *)x, (const float *)y, (float *)out);
    case ACCBLAS_F64: return dot_typed<double>(s, false, n, (const double *)x, (const double *)y, (double *)out);
    case ACCBLAS_C32: return dot_typed<float2>(s, conj_x != 0, n, (const float2 *)x, (const float2 *)y, (float2 *)out);
    default:          return dot_typed<double2>(s, conj_x != 0, n, (const double2 *)x, (const double2 *)y, (double2 *)out);
    }
}

}  // namespace accblas
