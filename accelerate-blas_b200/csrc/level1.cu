// level1.cu -- the remaining Level-1 reductions of the reference, each ONE pass over HBM:
//
//   sdot z xs ys   BLAS/Level1.hs:44-62    out = z + sum_i x_i * y_i, products and sum in DOUBLE
//                                          precision whatever the element type, rounded once
//   asum xs        BLAS/Level1.hs:100-109  out = sum_i |x_i|            (complex: |re| + |im|)
//   amax xs        BLAS/Level1.hs:115-126  index of the element of largest  |x_i| (same magnitude)
//   amin xs        BLAS/Level1.hs:132-143  index of the element of smallest |x_i|
//
// The reference evaluates these as Accelerate folds (no BLAS call, no foreign hook); these are NEW
// hooks with the same skeleton as dot.cu: a grid of <= 2 CTAs per SM streaming 16-byte vectors
// (ld.global.nc.L1::no_allocate, 4 loads in flight per thread per operand), register accumulation,
// shuffle + shared-memory block reduce, one partial per CTA, a one-CTA second pass combining the
// partials in index order.  A vector short enough for one CTA is finished by that CTA (one launch).
//
// amax / amin tie rule.  The reference folds with `cmp ix iy = snd ix > snd iy ? ix : iy`
// (`<` for amin): on equal magnitudes the RIGHT operand wins, so under any order-preserving
// association the LAST index among equal extrema is returned (cblas_i?amax returns the first).
// The combine used here -- larger magnitude, then larger index -- is associative and commutative and
// yields exactly that index; magnitudes are computed in the element's own precision like the
// reference (`abs (real c) + abs (imag c)` in Float for Complex Float), so the index is bit-exact.
#include "common.cuh"

namespace accblas {
namespace {

constexpr int L1_THREADS = 512;
constexpr int L1_UNROLL = 4;

template <typename T> union Vec16 {
    int4 raw;
    T e[16 / sizeof(T)];
};

__device__ __forceinline__ double2 zero2() { return make_double2(0.0, 0.0); }

// ---- double-precision products -------------------------------------------------------------
__device__ __forceinline__ void dfma(double &acc, float a, float b) { acc = fma((double)a, (double)b, acc); }
__device__ __forceinline__ void dfma(double &acc, double a, double b) { acc = fma(a, b, acc); }
__device__ __forceinline__ void dfma(double2 &acc, float2 a, float2 b) {
    const double ar = a.x, ai = a.y, br = b.x, bi = b.y;
    acc.x = fma(ar, br, acc.x); acc.x = fma(-ai, bi, acc.x);
    acc.y = fma(ar, bi, acc.y); acc.y = fma(ai, br, acc.y);
}
__device__ __forceinline__ void dfma(double2 &acc, double2 a, double2 b) {
    acc.x = fma(a.x, b.x, acc.x); acc.x = fma(-a.y, b.y, acc.x);
    acc.y = fma(a.x, b.y, acc.y); acc.y = fma(a.y, b.x, acc.y);
}
template <typename T> struct Wide { using type = double; };
template <> struct Wide<float2> { using type = double2; };
template <> struct Wide<double2> { using type = double2; };
template <typename W> __device__ __forceinline__ W wzero();
template <> __device__ __forceinline__ double wzero<double>() { return 0.0; }
template <> __device__ __forceinline__ double2 wzero<double2>() { return zero2(); }

__device__ __forceinline__ float  narrow_to(float *, double v) { return (float)v; }
__device__ __forceinline__ double narrow_to(double *, double v) { return v; }
__device__ __forceinline__ float2 narrow_to(float2 *, double2 v) { return make_float2((float)v.x, (float)v.y); }
__device__ __forceinline__ double2 narrow_to(double2 *, double2 v) { return v; }

// ---- magnitudes in the element's own precision (BLAS/Level1.hs:108,125,142) -----------------
__device__ __forceinline__ float  mag(float v) { return fabsf(v); }
__device__ __forceinline__ double mag(double v) { return fabs(v); }
__device__ __forceinline__ float  mag(float2 v) { return __fadd_rn(fabsf(v.x), fabsf(v.y)); }
__device__ __forceinline__ double mag(double2 v) { return __dadd_rn(fabs(v.x), fabs(v.y)); }

template <typename W>
__device__ __forceinline__ W block_sum(W v, W *smem)
{
    v = warp_sum(v);
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    if (lane == 0) smem[warp] = v;
    __syncthreads();
    if (warp == 0) {
        v = lane < (int)(blockDim.x >> 5) ? smem[lane] : wzero<W>();
        v = warp_sum(v);
    }
    return v;   // valid in warp 0
}

// ---- streaming traversals: f(slot, index, x_i [, y_i]); slot < L1_UNROLL names an independent chain ----
template <typename T, bool ALIGNED, typename F>
__device__ __forceinline__ void for_each1(int64_t n, const T *__restrict__ x, F f)
{
    constexpr int VEC = 16 / sizeof(T);
    if (ALIGNED) {
        const int64_t nv = n / VEC;
        constexpr int64_t CHUNK = (int64_t)L1_THREADS * L1_UNROLL;
        const int64_t nchunks = nv / CHUNK;
        const int4 *xv = reinterpret_cast<const int4 *>(x);
        for (int64_t c = blockIdx.x; c < nchunks; c += gridDim.x) {
            const int64_t base = c * CHUNK + threadIdx.x;
            Vec16<T> a[L1_UNROLL];
#pragma unroll
            for (int u = 0; u < L1_UNROLL; ++u) a[u].raw = ldg_stream(xv + base + u * L1_THREADS);
#pragma unroll
            for (int u = 0; u < L1_UNROLL; ++u)
#pragma unroll
                for (int e = 0; e < VEC; ++e) f(u, (base + u * L1_THREADS) * VEC + e, a[u].e[e]);
        }
        for (int64_t v = nchunks * CHUNK + (int64_t)blockIdx.x * L1_THREADS + threadIdx.x; v < nv;
             v += (int64_t)gridDim.x * L1_THREADS) {
            Vec16<T> a;
            a.raw = ldg_stream(xv + v);
#pragma unroll
            for (int e = 0; e < VEC; ++e) f(0, v * VEC + e, a.e[e]);
        }
        if (blockIdx.x == 0 && threadIdx.x < (int)(n - nv * VEC)) {
            const int64_t i = nv * VEC + threadIdx.x;
            f(1, i, x[i]);
        }
    } else {
        for (int64_t i = (int64_t)blockIdx.x * L1_THREADS + threadIdx.x; i < n; i += (int64_t)gridDim.x * L1_THREADS)
            f(0, i, x[i]);
    }
}

template <typename T, bool ALIGNED, typename F>
__device__ __forceinline__ void for_each2(int64_t n, const T *__restrict__ x, const T *__restrict__ y, F f)
{
    constexpr int VEC = 16 / sizeof(T);
    if (ALIGNED) {
        const int64_t nv = n / VEC;
        constexpr int64_t CHUNK = (int64_t)L1_THREADS * L1_UNROLL;
        const int64_t nchunks = nv / CHUNK;
        const int4 *xv = reinterpret_cast<const int4 *>(x);
        const int4 *yv = reinterpret_cast<const int4 *>(y);
        for (int64_t c = blockIdx.x; c < nchunks; c += gridDim.x) {
            const int64_t base = c * CHUNK + threadIdx.x;
            Vec16<T> a[L1_UNROLL], b[L1_UNROLL];
#pragma unroll
            for (int u = 0; u < L1_UNROLL; ++u) a[u].raw = ldg_stream(xv + base + u * L1_THREADS);
#pragma unroll
            for (int u = 0; u < L1_UNROLL; ++u) b[u].raw = ldg_stream(yv + base + u * L1_THREADS);
#pragma unroll
            for (int u = 0; u < L1_UNROLL; ++u)
#pragma unroll
                for (int e = 0; e < VEC; ++e) f(u, a[u].e[e], b[u].e[e]);
        }
        for (int64_t v = nchunks * CHUNK + (int64_t)blockIdx.x * L1_THREADS + threadIdx.x; v < nv;
             v += (int64_t)gridDim.x * L1_THREADS) {
            Vec16<T> a, b;
            a.raw = ldg_stream(xv + v);
            b.raw = ldg_stream(yv + v);
#pragma unroll
            for (int e = 0; e < VEC; ++e) f(0, a.e[e], b.e[e]);
        }
        if (blockIdx.x == 0 && threadIdx.x < (int)(n - nv * VEC)) {
            const int64_t i = nv * VEC + threadIdx.x;
            f(1, x[i], y[i]);
        }
    } else {
        for (int64_t i = (int64_t)blockIdx.x * L1_THREADS + threadIdx.x; i < n; i += (int64_t)gridDim.x * L1_THREADS)
            f(0, x[i], y[i]);
    }
}

// =============================================================================================
// sdot
// =============================================================================================
template <typename T, bool ALIGNED>
__global__ void __launch_bounds__(L1_THREADS, 2)
sdot_partial_kernel(int64_t n, const T *__restrict__ x, const T *__restrict__ y, typename Wide<T>::type z,
                    typename Wide<T>::type *__restrict__ partial, T *__restrict__ out)
{
    using W = typename Wide<T>::type;
    __shared__ W red[L1_THREADS / 32];
    W acc[L1_UNROLL];
#pragma unroll
    for (int u = 0; u < L1_UNROLL; ++u) acc[u] = wzero<W>();
    for_each2<T, ALIGNED>(n, x, y, [&](int u, T a, T b) { dfma(acc[u], a, b); });
    W s = acc[0];
#pragma unroll
    for (int u = 1; u < L1_UNROLL; ++u) s = add_s(s, acc[u]);
    s = block_sum(s, red);
    if (threadIdx.x == 0) {
        if (gridDim.x == 1) *out = narrow_to(out, add_s(z, s));      // short vector: finished here
        else partial[blockIdx.x] = s;
    }
}

template <typename T>
__global__ void __launch_bounds__(L1_THREADS)
sdot_final_kernel(int nparts, const typename Wide<T>::type *__restrict__ partial, typename Wide<T>::type z,
                  T *__restrict__ out)
{
    using W = typename Wide<T>::type;
    __shared__ W red[L1_THREADS / 32];
    W s = wzero<W>();
    for (int i = threadIdx.x; i < nparts; i += L1_THREADS) s = add_s(s, partial[i]);
    s = block_sum(s, red);
    if (threadIdx.x == 0) *out = narrow_to(out, add_s(z, s));
}

template <typename T> bool aligned16(const T *p) { return (reinterpret_cast<uintptr_t>(p) & 15u) == 0; }

// CTAs for an n-element pass: one per L1_THREADS vectors of work, at most 2 per SM
template <typename T>
int grid_for(int64_t n, bool aligned)
{
    constexpr int VEC = 16 / sizeof(T);
    const int64_t work = aligned ? (n / VEC + L1_THREADS - 1) / L1_THREADS : (n + L1_THREADS - 1) / L1_THREADS;
    const int cap = 2 * sm_count();
    // up to 8 vectors per thread stay in one CTA: one launch beats two for short vectors
    if (work <= 8) return 1;
    return (int)(work < cap ? work : cap);
}

template <typename T>
int sdot_typed(cudaStream_t st, int64_t n, const void *z_host, const T *x, const T *y, T *out)
{
    using W = typename Wide<T>::type;
    W z;
    if constexpr (Scalar<T>::cplx) {
        const T zz = *static_cast<const T *>(z_host);
        z = make_double2((double)zz.x, (double)zz.y);
    } else {
        z = (double)*static_cast<const T *>(z_host);
    }
    if (n == 0) {                                  // fold (+) z over nothing = z
        ACC_CUDA(cudaMemcpyAsync(out, z_host, sizeof(T), cudaMemcpyHostToDevice, st));
        return ACCBLAS_OK;
    }
    const bool al = aligned16(x) && aligned16(y);
    const int grid = grid_for<T>(n, al);
    W *partial = nullptr;
    if (grid > 1) ACC_CUDA(scratch_alloc(reinterpret_cast<void **>(&partial), sizeof(W) * (size_t)grid, st));
    if (al) sdot_partial_kernel<T, true><<<grid, L1_THREADS, 0, st>>>(n, x, y, z, partial, out);
    else    sdot_partial_kernel<T, false><<<grid, L1_THREADS, 0, st>>>(n, x, y, z, partial, out);
    int s = check_launch("sdot_partial_kernel");
    if (grid > 1) {
        if (s == ACCBLAS_OK) {
            sdot_final_kernel<T><<<1, L1_THREADS, 0, st>>>(grid, partial, z, out);
            s = check_launch("sdot_final_kernel");
        }
        cudaError_t e = cudaFreeAsync(partial, st);
        if (s == ACCBLAS_OK && e != cudaSuccess) return cuda_fail(e, "cudaFreeAsync");
    }
    return s;
}

// =============================================================================================
// asum: sum of magnitudes, accumulated in double, rounded once to the element's base type
// =============================================================================================
template <typename T, bool ALIGNED>
__global__ void __launch_bounds__(L1_THREADS, 2)
asum_partial_kernel(int64_t n, const T *__restrict__ x, double *__restrict__ partial,
                    typename Scalar<T>::real *__restrict__ out)
{
    __shared__ double red[L1_THREADS / 32];
    double acc[L1_UNROLL];
#pragma unroll
    for (int u = 0; u < L1_UNROLL; ++u) acc[u] = 0.0;
    for_each1<T, ALIGNED>(n, x, [&](int u, int64_t, T a) {
        if constexpr (Scalar<T>::cplx) acc[u] += (double)fabs(a.x) + (double)fabs(a.y);
        else acc[u] += (double)fabs(a);
    });
    double s = (acc[0] + acc[1]) + (acc[2] + acc[3]);
    s = block_sum(s, red);
    if (threadIdx.x == 0) {
        if (gridDim.x == 1) *out = (typename Scalar<T>::real)s;
        else partial[blockIdx.x] = s;
    }
}

template <typename R>
__global__ void __launch_bounds__(L1_THREADS)
asum_final_kernel(int nparts, const double *__restrict__ partial, R *__restrict__ out)
{
    __shared__ double red[L1_THREADS / 32];
    double s = 0.0;
    for (int i = threadIdx.x; i < nparts; i += L1_THREADS) s += partial[i];
    s = block_sum(s, red);
    if (threadIdx.x == 0) *out = (R)s;
}

template <typename T>
int asum_typed(cudaStream_t st, int64_t n, const T *x, void *out_v)
{
    using R = typename Scalar<T>::real;
    R *out = static_cast<R *>(out_v);
    if (n == 0) {                                  // sum of nothing
        ACC_CUDA(cudaMemsetAsync(out, 0, sizeof(R), st));
        return ACCBLAS_OK;
    }
    const bool al = aligned16(x);
    const int grid = grid_for<T>(n, al);
    double *partial = nullptr;
    if (grid > 1) ACC_CUDA(scratch_alloc(reinterpret_cast<void **>(&partial), sizeof(double) * (size_t)grid, st));
    if (al) asum_partial_kernel<T, true><<<grid, L1_THREADS, 0, st>>>(n, x, partial, out);
    else    asum_partial_kernel<T, false><<<grid, L1_THREADS, 0, st>>>(n, x, partial, out);
    int s = check_launch("asum_partial_kernel");
    if (grid > 1) {
        if (s == ACCBLAS_OK) {
            asum_final_kernel<R><<<1, L1_THREADS, 0, st>>>(grid, partial, out);
            s = check_launch("asum_final_kernel");
        }
        cudaError_t e = cudaFreeAsync(partial, st);
        if (s == ACCBLAS_OK && e != cudaSuccess) return cuda_fail(e, "cudaFreeAsync");
    }
    return s;
}

// =============================================================================================
// amax / amin: (magnitude, index) pairs; combine = better magnitude, then LARGER index
// =============================================================================================
template <typename R> struct Best { R v; int64_t i; };

template <bool MAX, typename R>
__device__ __forceinline__ bool better(R v, int64_t i, R bv, int64_t bi)
{
    // "better or equal-and-later": the reference's cmp keeps the right operand unless the left is
    // strictly better (BLAS/Level1.hs:124,141).  bi < 0 marks an empty slot.
    if (bi < 0) return true;
    if (i < 0) return false;
    return MAX ? (v > bv || (v == bv && i > bi)) : (v < bv || (v == bv && i > bi));
}

template <bool MAX, typename R>
__device__ __forceinline__ Best<R> warp_best(Best<R> b)
{
#pragma unroll
    for (int m = 16; m > 0; m >>= 1) {
        const R ov = __shfl_xor_sync(0xffffffffu, b.v, m);
        const int64_t oi = __shfl_xor_sync(0xffffffffu, b.i, m);
        if (better<MAX>(ov, oi, b.v, b.i)) { b.v = ov; b.i = oi; }
    }
    return b;
}

template <bool MAX, typename R>
__device__ __forceinline__ Best<R> block_best(Best<R> b, R *sv, int64_t *si)
{
    b = warp_best<MAX>(b);
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    if (lane == 0) { sv[warp] = b.v; si[warp] = b.i; }
    __syncthreads();
    if (warp == 0) {
        if (lane < (int)(blockDim.x >> 5)) { b.v = sv[lane]; b.i = si[lane]; }
        else { b.v = (R)0; b.i = -1; }
        b = warp_best<MAX>(b);
    }
    return b;   // valid in warp 0
}

template <typename T, bool MAX, bool ALIGNED>
__global__ void __launch_bounds__(L1_THREADS, 2)
iamax_partial_kernel(int64_t n, const T *__restrict__ x, typename Scalar<T>::real *__restrict__ pv,
                     int64_t *__restrict__ pi, int64_t *__restrict__ out)
{
    using R = typename Scalar<T>::real;
    __shared__ R sv[L1_THREADS / 32];
    __shared__ int64_t si[L1_THREADS / 32];
    Best<R> b[L1_UNROLL];
#pragma unroll
    // Within a slot the indices arrive in increasing order, so "equal magnitude: later index wins" is just
    // >= (<= for amin) against the running best: one compare per element, the 64-bit index is only written on
    // an update.  Magnitudes are >= 0, so -1 / +inf make the first element of a slot win; a slot that saw
    // nothing (or only NaNs, which never compare true) keeps index -1 and loses every combine below.
    for (int u = 0; u < L1_UNROLL; ++u) { b[u].v = MAX ? (R)-1 : (R)INFINITY; b[u].i = -1; }
    for_each1<T, ALIGNED>(n, x, [&](int u, int64_t i, T a) {
        const R v = mag(a);
        if (MAX ? v >= b[u].v : v <= b[u].v) { b[u].v = v; b[u].i = i; }
    });
#pragma unroll
    for (int u = 1; u < L1_UNROLL; ++u)
        if (better<MAX>(b[u].v, b[u].i, b[0].v, b[0].i)) b[0] = b[u];
    Best<R> r = block_best<MAX>(b[0], sv, si);
    if (threadIdx.x == 0) {
        if (gridDim.x == 1) *out = r.i < 0 ? n - 1 : r.i;   // all NaN: the reference's fold ends on the last index
        else { pv[blockIdx.x] = r.v; pi[blockIdx.x] = r.i; }
    }
}

template <typename R, bool MAX>
__global__ void __launch_bounds__(L1_THREADS)
iamax_final_kernel(int nparts, int64_t n, const R *__restrict__ pv, const int64_t *__restrict__ pi, int64_t *__restrict__ out)
{
    __shared__ R sv[L1_THREADS / 32];
    __shared__ int64_t si[L1_THREADS / 32];
    Best<R> b; b.v = (R)0; b.i = -1;
    for (int i = threadIdx.x; i < nparts; i += L1_THREADS)
        if (better<MAX>(pv[i], pi[i], b.v, b.i)) { b.v = pv[i]; b.i = pi[i]; }
    b = block_best<MAX>(b, sv, si);
    if (threadIdx.x == 0) *out = b.i < 0 ? n - 1 : b.i;
}

template <typename T, bool MAX>
int iamax_typed(cudaStream_t st, int64_t n, const T *x, int64_t *out)
{
    using R = typename Scalar<T>::real;
    const bool al = aligned16(x);
    const int grid = grid_for<T>(n, al);
    char *ws = nullptr;
    R *pv = nullptr; int64_t *pi = nullptr;
    if (grid > 1) {
        ACC_CUDA(scratch_alloc(reinterpret_cast<void **>(&ws), 16 * (size_t)grid, st));
        pi = reinterpret_cast<int64_t *>(ws);
        pv = reinterpret_cast<R *>(ws + 8 * (size_t)grid);
    }
    if (al) iamax_partial_kernel<T, MAX, true><<<grid, L1_THREADS, 0, st>>>(n, x, pv, pi, out);
    else    iamax_partial_kernel<T, MAX, false><<<grid, L1_THREADS, 0, st>>>(n, x, pv, pi, out);
    int s = check_launch("iamax_partial_kernel");
    if (grid > 1) {
        if (s == ACCBLAS_OK) {
            iamax_final_kernel<R, MAX><<<1, L1_THREADS, 0, st>>>(grid, n, pv, pi, out);
            s = check_launch("iamax_final_kernel");
        }
        cudaError_t e = cudaFreeAsync(ws, st);
        if (s == ACCBLAS_OK && e != cudaSuccess) return cuda_fail(e, "cudaFreeAsync");
    }
    return s;
}

}  // namespace

int sdot_dispatch(cudaStream_t s, int dtype, int64_t n, const void *z_host, const void *x, const void *y, void *out)
{
    switch (dtype) {
    case ACCBLAS_F32: return sdot_typed<float>(s, n, z_host, (const float *)x, (const float *)y, (float *)out);
    case ACCBLAS_F64: return sdot_typed<double>(s, n, z_host, (const double *)x, (const double *)y, (double *)out);
    case ACCBLAS_C32: return sdot_typed<float2>(s, n, z_host, (const float2 *)x, (const float2 *)y, (float2 *)out);
    default:          return sdot_typed<double2>(s, n, z_host, (const double2 *)x, (const double2 *)y, (double2 *)out);
    }
}

int asum_dispatch(cudaStream_t s, int dtype, int64_t n, const void *x, void *out)
{
    switch (dtype) {
    case ACCBLAS_F32: return asum_typed<float>(s, n, (const float *)x, out);
    case ACCBLAS_F64: return asum_typed<double>(s, n, (const double *)x, out);
    case ACCBLAS_C32: return asum_typed<float2>(s, n, (const float2 *)x, out);
    default:          return asum_typed<double2>(s, n, (const double2 *)x, out);
    }
}

int iamax_dispatch(cudaStream_t s, int dtype, int want_max, int64_t n, const void *x, int64_t *out)
{
#define GO(T) (want_max ? iamax_typed<T, true>(s, n, (const T *)x, out) : iamax_typed<T, false>(s, n, (const T *)x, out))
    switch (dtype) {
    case ACCBLAS_F32: return GO(float);
    case ACCBLAS_F64: return GO(double);
    case ACCBLAS_C32: return GO(float2);
    default:          return GO(double2);
    }
#undef GO
}

}  // namespace accblas
