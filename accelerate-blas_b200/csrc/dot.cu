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
//
// Two more users of the same skeleton:
//  * the COMPENSATED dot (accblas_dot_compensated; SURVEY 8(f)-2, the reference's Numeric/Sum.hs): every product is
//    split exactly (p, e) with one FMA, the running sum is a TwoSum whose rounding errors are collected beside it
//    ("Dot2" of Ogita, Rump and Oishi), and partial sums of threads, warps and CTAs are merged with the reference's
//    own Kahan-Babuska-Neumaier combine (Numeric/Sum.hs:119-125, kbnAdd) -- the result is as accurate as a dot
//    computed in twice the working precision and rounded once.  Float operands are multiplied and summed in double.
//  * the SHARDED dot (accblas_mg_dot): one launch; the last CTA to finish sums the partials in index order and runs
//    the scalar all-reduce over the peers' exchange buffers itself (mg_exchange.cuh).
#include "common.cuh"
#include "mg_exchange.cuh"

namespace accblas {
namespace {

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

// ----------------------------------------------------------------------------------------------------------------
// accumulation policies
// ----------------------------------------------------------------------------------------------------------------
// Plain: per-thread sums in the element type (the reference's fold type), cross-thread stages in double.
template <typename T, int UNROLL> struct Plain {
    static constexpr int SLOTS = UNROLL;
    using Slot = T;
    using Part = typename Acc<T>::type;
    __device__ static __forceinline__ Slot zero() { return Scalar<T>::zero(); }
    __device__ static __forceinline__ void step(Slot &s, T a, T b) { s = fma_acc(a, b, s); }
    __device__ static __forceinline__ Part lift(Slot s) { return widen(s); }
    __device__ static __forceinline__ Part pzero() {
        if constexpr (sizeof(Part) == 16) return make_double2(0.0, 0.0); else return 0.0;
    }
    __device__ static __forceinline__ Part add(Part a, Part b) { return add_s(a, b); }
    __device__ static __forceinline__ Part shfl(Part v, int m) { return shfl_xor_s(v, m); }
    __device__ static __forceinline__ T finish(Part p) { return narrow<T>(p); }
};

// Compensated: (sum, compensation) per real component, all in double.  The intrinsics keep nvcc from contracting
// a product into the following add -- the error terms are only exact for the operations as written.
template <int N> struct Kbn { double s[N], c[N]; };

__device__ __forceinline__ void kbn_push(double &s, double &c, double a, double b, bool exact_product)
{
    const double p = __dmul_rn(a, b);
    const double e = exact_product ? 0.0 : __fma_rn(a, b, -p);            // a * b = p + e exactly
    const double t = __dadd_rn(s, p);                                      // TwoSum: s + p = t + err exactly
    const double z = __dsub_rn(t, s);
    const double err = __dadd_rn(__dsub_rn(s, __dsub_rn(t, z)), __dsub_rn(p, z));
    s = t;
    c = __dadd_rn(c, __dadd_rn(err, e));
}
// the reference's kbnAdd (Numeric/Sum.hs:119-125): s' = s1 + s2; c' = c1 + c2 + (the rounding error of s1 + s2,
// recovered from the larger operand)
__device__ __forceinline__ void kbn_merge(double &s1, double &c1, double s2, double c2)
{
    const double s = __dadd_rn(s1, s2);
    const double lost = fabs(s1) >= fabs(s2) ? __dadd_rn(__dsub_rn(s1, s), s2) : __dadd_rn(__dsub_rn(s2, s), s1);
    c1 = __dadd_rn(__dadd_rn(c1, c2), lost);
    s1 = s;
}

template <typename T, int UNROLL> struct Comp {
    static constexpr int SLOTS = 1;                                        // the chain is short next to the loads
    static constexpr int N = Scalar<T>::cplx ? 2 : 1;
    static constexpr bool EXACT = sizeof(typename Scalar<T>::real) == 4;   // float x float is exact in double
    using Slot = Kbn<N>;
    using Part = Kbn<N>;
    __device__ static __forceinline__ Slot zero() {
        Slot z;
#pragma unroll
        for (int i = 0; i < N; ++i) { z.s[i] = 0.0; z.c[i] = 0.0; }
        return z;
    }
    __device__ static __forceinline__ void step(Slot &k, T a, T b) {
        if constexpr (N == 1) {
            kbn_push(k.s[0], k.c[0], (double)a, (double)b, EXACT);
        } else {
            kbn_push(k.s[0], k.c[0], (double)a.x, (double)b.x, EXACT);
            kbn_push(k.s[0], k.c[0], -(double)a.y, (double)b.y, EXACT);
            kbn_push(k.s[1], k.c[1], (double)a.x, (double)b.y, EXACT);
            kbn_push(k.s[1], k.c[1], (double)a.y, (double)b.x, EXACT);
        }
    }
    __device__ static __forceinline__ Part lift(Slot s) { return s; }
    __device__ static __forceinline__ Part pzero() { return zero(); }
    __device__ static __forceinline__ Part add(Part a, Part b) {
#pragma unroll
        for (int i = 0; i < N; ++i) kbn_merge(a.s[i], a.c[i], b.s[i], b.c[i]);
        return a;
    }
    __device__ static __forceinline__ Part shfl(Part v, int m) {
        Part r;
#pragma unroll
        for (int i = 0; i < N; ++i) {
            r.s[i] = __shfl_xor_sync(0xffffffffu, v.s[i], m);
            r.c[i] = __shfl_xor_sync(0xffffffffu, v.c[i], m);
        }
        return r;
    }
    __device__ static __forceinline__ T finish(Part p) {
        if constexpr (N == 1) return narrow<T>(__dadd_rn(p.s[0], p.c[0]));
        else return narrow<T>(make_double2(__dadd_rn(p.s[0], p.c[0]), __dadd_rn(p.s[1], p.c[1])));
    }
};

template <typename P>
__device__ __forceinline__ typename P::Part block_sum(typename P::Part v, typename P::Part *smem)
{
#pragma unroll
    for (int m = 16; m > 0; m >>= 1) v = P::add(v, P::shfl(v, m));
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    if (lane == 0) smem[warp] = v;
    __syncthreads();
    if (warp == 0) {
        v = lane < (int)(blockDim.x >> 5) ? smem[lane] : P::pzero();
#pragma unroll
        for (int m = 16; m > 0; m >>= 1) v = P::add(v, P::shfl(v, m));
    }
    return v;   // valid in warp 0
}

// One CTA's share of the products, reduced to one Part (valid in warp 0).
template <typename T, typename P, bool CONJ, bool ALIGNED, int THREADS, int UNROLL>
__device__ __forceinline__ typename P::Part dot_cta(int64_t n, const T *__restrict__ x, const T *__restrict__ y,
                                                    typename P::Part *red)
{
    constexpr int VEC = 16 / sizeof(T);
    typename P::Slot acc[P::SLOTS];
#pragma unroll
    for (int u = 0; u < P::SLOTS; ++u) acc[u] = P::zero();

    if (ALIGNED) {
        const int64_t nv = n / VEC;
        constexpr int64_t CHUNK = (int64_t)THREADS * UNROLL;
        const int64_t nchunks = nv / CHUNK;
        const int4 *xv = reinterpret_cast<const int4 *>(x);
        const int4 *yv = reinterpret_cast<const int4 *>(y);
        for (int64_t c = blockIdx.x; c < nchunks; c += gridDim.x) {
            const int64_t base = c * CHUNK + threadIdx.x;
            Vec16<T> a[UNROLL], b[UNROLL];
#pragma unroll
            for (int u = 0; u < UNROLL; ++u) a[u].raw = ldg_stream(xv + base + u * THREADS);
#pragma unroll
            for (int u = 0; u < UNROLL; ++u) b[u].raw = ldg_stream(yv + base + u * THREADS);
#pragma unroll
            for (int u = 0; u < UNROLL; ++u)
#pragma unroll
                for (int e = 0; e < VEC; ++e)
                    P::step(acc[u % P::SLOTS], CONJ ? conj_s(a[u].e[e]) : a[u].e[e], b[u].e[e]);
        }
        // remainder: whole vectors past the last full chunk, then the scalar tail
        for (int64_t v = nchunks * CHUNK + (int64_t)blockIdx.x * THREADS + threadIdx.x; v < nv;
             v += (int64_t)gridDim.x * THREADS) {
            Vec16<T> a, b;
            a.raw = ldg_stream(xv + v);
            b.raw = ldg_stream(yv + v);
#pragma unroll
            for (int e = 0; e < VEC; ++e) P::step(acc[0], CONJ ? conj_s(a.e[e]) : a.e[e], b.e[e]);
        }
        if (blockIdx.x == 0 && threadIdx.x < (int)(n - nv * VEC)) {
            const int64_t i = nv * VEC + threadIdx.x;
            P::step(acc[P::SLOTS > 1 ? 1 : 0], CONJ ? conj_s(x[i]) : x[i], y[i]);
        }
    } else {
        for (int64_t i = (int64_t)blockIdx.x * THREADS + threadIdx.x; i < n; i += (int64_t)gridDim.x * THREADS)
            P::step(acc[0], CONJ ? conj_s(x[i]) : x[i], y[i]);
    }
    typename P::Part s = P::lift(acc[0]);
#pragma unroll
    for (int u = 1; u < P::SLOTS; ++u) s = P::add(s, P::lift(acc[u]));
    return block_sum<P>(s, red);
}

template <typename T, typename P, bool CONJ, bool ALIGNED, int THREADS, int UNROLL, int MINB>
__global__ void __launch_bounds__(THREADS, MINB)
dot_partial_kernel(int64_t n, const T *__restrict__ x, const T *__restrict__ y, typename P::Part *__restrict__ partial)
{
    __shared__ typename P::Part red[THREADS / 32];
    const typename P::Part s = dot_cta<T, P, CONJ, ALIGNED, THREADS, UNROLL>(n, x, y, red);
    if (threadIdx.x == 0) partial[blockIdx.x] = s;
}

template <typename T, typename P, int THREADS>
__global__ void __launch_bounds__(THREADS)
dot_final_kernel(int nparts, const typename P::Part *__restrict__ partial, T *__restrict__ out)
{
    __shared__ typename P::Part red[THREADS / 32];
    typename P::Part s = P::pzero();
    for (int i = threadIdx.x; i < nparts; i += THREADS) s = P::add(s, partial[i]);
    s = block_sum<P>(s, red);
    if (threadIdx.x == 0) *out = P::finish(s);
}

// Sharded dot in one launch: partials as above; the LAST CTA to arrive (atomicInc wraps the counter back to zero for
// the next call) sums them in index order -- the same tree as dot_final_kernel -- and then is this rank's side of the
// scalar all-reduce: thread r deposits the value in rank r's exchange buffer, raises this rank's flag there, waits for
// rank r's deposit here; thread 0 adds the deposits in rank order (every rank computes the same bits).
template <typename T, typename P, bool CONJ, bool ALIGNED, int THREADS, int UNROLL, int MINB>
__global__ void __launch_bounds__(THREADS, MINB)
dot_mg_kernel(int64_t n, const T *__restrict__ x, const T *__restrict__ y, T *__restrict__ out, DotExchange ex)
{
    using Part = typename P::Part;
    __shared__ Part red[THREADS / 32];
    __shared__ T got[MG_MAX_RANKS];
    __shared__ T mine;
    __shared__ int is_last;
    Part *partial = static_cast<Part *>(ex.partials);
    Part s = dot_cta<T, P, CONJ, ALIGNED, THREADS, UNROLL>(n, x, y, red);
    if (threadIdx.x == 0) {
        partial[blockIdx.x] = s;
        __threadfence();
        is_last = atomicInc(ex.counter, gridDim.x - 1) == gridDim.x - 1;
    }
    __syncthreads();
    if (!is_last) return;
    __threadfence();
    s = P::pzero();
    for (int i = threadIdx.x; i < (int)gridDim.x; i += THREADS) {
        Part v;
        const double *src = reinterpret_cast<const double *>(partial + i);
        double *dst = reinterpret_cast<double *>(&v);
#pragma unroll
        for (int w = 0; w < (int)(sizeof(Part) / 8); ++w) dst[w] = __ldcv(src + w);
        s = P::add(s, v);
    }
    s = block_sum<P>(s, red);
    if (threadIdx.x == 0) mine = P::finish(s);
    __syncthreads();
    const int r = threadIdx.x;
    if (ex.world == 0) {                              // one GPU (accblas_dot): the last CTA's sum is the result
        if (r == 0) *out = mine;
        return;
    }
    if (r < ex.world) {
        char *dst = ex.peers.p[r] + ex.parity_off;
        *reinterpret_cast<T *>(dst + XFLAGS + (size_t)ex.rank * XSLOT) = mine;
        __threadfence_system();
        st_release_sys(reinterpret_cast<uint32_t *>(dst + 64 * ex.rank), ex.seq);
        const char *own = ex.peers.p[ex.rank] + ex.parity_off;
        wait_flag(reinterpret_cast<const uint32_t *>(own + 64 * r), ex.seq);
        got[r] = __ldcv(reinterpret_cast<const T *>(own + XFLAGS + (size_t)r * XSLOT));
    }
    __syncthreads();
    if (r == 0) {
        T acc = got[0];
        for (int i = 1; i < ex.world; ++i) acc = add_s(acc, got[i]);
        *out = acc;
    }
}

template <typename T> inline bool both_aligned(const T *x, const T *y) {
    return ((reinterpret_cast<uintptr_t>(x) | reinterpret_cast<uintptr_t>(y)) & 15u) == 0;
}
template <typename T> inline int grid_for(int64_t n, bool aligned, int threads, int per_sm) {
    constexpr int VEC = 16 / sizeof(T);
    const int64_t work = aligned ? (n / VEC + threads - 1) / threads : (n + threads - 1) / threads;
    int grid = per_sm * sm_count();
    if (work < grid) grid = (int)(work > 0 ? work : 1);
    return grid;
}

// ONE launch for the single-GPU dot as well: the last CTA to arrive sums the partials (dot_mg_kernel with world = 0).  That
// needs an arrival counter that is zero at launch and a partial buffer -- 33 KiB of device memory kept per (thread, device,
// stream): calls on one stream are ordered, so they can share a counter that wraps back to zero by itself, and two
// streams never share one.  Up to 32 streams per thread; beyond that (or if the allocation fails, e.g. under stream
// capture) the two-kernel path runs.  A launch costs ~3 us on these GPUs: ddot 2^20 10.3 -> 8.2 us.
constexpr int DOT_SLOT_GRID = 1024;
char *dot_slot(cudaStream_t st)
{
    struct Entry { int dev; cudaStream_t st; char *mem; };
    // the slots die with the host thread that made them (cudaFree waits for the launches still using one; at process
    // exit the runtime may already be gone and the call fails harmlessly)
    struct Slots {
        Entry e[32];
        int count = 0;
        ~Slots() { for (int i = 0; i < count; ++i) cudaFree(e[i].mem); }
    };
    static thread_local Slots slots;
    Entry *entries = slots.e;
    int &count = slots.count;
    if (g_tuning[ACCBLAS_TUNE_DOT_FUSED] == 0) return nullptr;
    // A capture must not bake a slot into a graph: the graph could be replayed on another stream, twice at once, or
    // beside eager dots of this stream, and all of those would share the arrival counter.  Checked BEFORE the cache
    // look-up -- a stream that already owns a slot can start capturing later.  Captured dots take the two-kernel path.
    cudaStreamCaptureStatus cap = cudaStreamCaptureStatusNone;
    if (cudaStreamIsCapturing(st, &cap) != cudaSuccess || cap != cudaStreamCaptureStatusNone) { cudaGetLastError(); return nullptr; }
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess) return nullptr;
    for (int i = 0; i < count; ++i)
        if (entries[i].dev == dev && entries[i].st == st) return entries[i].mem;
    if (count == 32) return nullptr;
    char *mem = nullptr;
    const size_t bytes = 64 + 32 * (size_t)DOT_SLOT_GRID;
    if (cudaMalloc(reinterpret_cast<void **>(&mem), bytes) != cudaSuccess || cudaMemset(mem, 0, bytes) != cudaSuccess) {
        cudaGetLastError();
        if (mem) cudaFree(mem);
        return nullptr;
    }
    entries[count++] = Entry{dev, st, mem};
    return mem;
}

template <typename T, typename P, int THREADS, int UNROLL, int MINB>
int dot_launch(cudaStream_t st, bool conj, int64_t n, const T *x, const T *y, T *out)
{
    using Part = typename P::Part;
    static_assert(sizeof(Part) <= 32, "a partial fits its 32-byte slot");
    const bool aligned = both_aligned(x, y);
    const int grid = grid_for<T>(n, aligned, THREADS, MINB);
    if (char *slot = grid <= DOT_SLOT_GRID ? dot_slot(st) : nullptr) {
        DotExchange ex{};
        ex.world = 0;
        ex.counter = reinterpret_cast<unsigned *>(slot);
        ex.partials = slot + 64;
        const bool cj = conj && Scalar<T>::cplx;
#define GO(CJ, AL) dot_mg_kernel<T, P, CJ, AL, THREADS, UNROLL, MINB><<<grid, THREADS, 0, st>>>(n, x, y, out, ex)
        if (aligned) { if (cj) GO(true, true); else GO(false, true); }
        else         { if (cj) GO(true, false); else GO(false, false); }
#undef GO
        return check_launch("dot_fused_kernel");
    }
    Part *partial = nullptr;
    ACC_CUDA(scratch_alloc(reinterpret_cast<void **>(&partial), sizeof(Part) * (size_t)grid, st));
    const bool cj = conj && Scalar<T>::cplx;
#define GO(CJ, AL) dot_partial_kernel<T, P, CJ, AL, THREADS, UNROLL, MINB><<<grid, THREADS, 0, st>>>(n, x, y, partial)
    if (aligned) { if (cj) GO(true, true); else GO(false, true); }
    else         { if (cj) GO(true, false); else GO(false, false); }
#undef GO
    int s = check_launch("dot_partial_kernel");
    if (s == ACCBLAS_OK) {
        dot_final_kernel<T, P, 512><<<1, 512, 0, st>>>(grid, partial, out);
        s = check_launch("dot_final_kernel");
    }
    cudaError_t e = cudaFreeAsync(partial, st);
    if (s == ACCBLAS_OK && e != cudaSuccess) return cuda_fail(e, "cudaFreeAsync");
    return s;
}

template <typename T>
int dot_typed(cudaStream_t st, bool conj, int64_t n, const T *x, const T *y, T *out, bool compensated)
{
    if (n == 0) {   // fold (+) 0 over nothing
        ACC_CUDA(cudaMemsetAsync(out, 0, sizeof(T), st));
        return ACCBLAS_OK;
    }
    if (compensated) return dot_launch<T, Comp<T, 4>, 512, 4, 2>(st, conj, n, x, y, out);
    switch (g_tuning[ACCBLAS_TUNE_DOT_SHAPE]) {                 // experiments only (profiles/r02_dot_shapes.jsonl)
    case 1:  return dot_launch<T, Plain<T, 4>, 256, 4, 4>(st, conj, n, x, y, out);
    case 2:  return dot_launch<T, Plain<T, 8>, 256, 8, 2>(st, conj, n, x, y, out);
    case 4:  return dot_launch<T, Plain<T, 8>, 512, 8, 1>(st, conj, n, x, y, out);
    case 5:  return dot_launch<T, Plain<T, 4>, 1024, 4, 1>(st, conj, n, x, y, out);
    case 0:  return dot_launch<T, Plain<T, 4>, 512, 4, 2>(st, conj, n, x, y, out);
    default:
        // Float: 8 loads per operand in flight per thread measure 7 056 GB/s at 2^28 against 6 639 for 4 (the 4-byte
        // elements need twice the FMAs per byte, and the deeper batch hides them); the 8-byte types are within 0.3 %
        // of each other on every shape (profiles/r02_dot_shapes.jsonl)
        if (sizeof(T) == 4) return dot_launch<T, Plain<T, 8>, 256, 8, 2>(st, conj, n, x, y, out);
        return dot_launch<T, Plain<T, 4>, 512, 4, 2>(st, conj, n, x, y, out);
    }
}

template <typename T>
int dot_mg_typed(cudaStream_t st, bool conj, int64_t n, const T *x, const T *y, T *out, const DotExchange &ex, int max_grid)
{
    using P = Plain<T, 4>;
    const bool aligned = both_aligned(x, y);
    int grid = grid_for<T>(n, aligned, 512, 2);
    if (grid > max_grid) grid = max_grid;
    const bool cj = conj && Scalar<T>::cplx;
#define GO(CJ, AL) dot_mg_kernel<T, P, CJ, AL, 512, 4, 2><<<grid, 512, 0, st>>>(n, x, y, out, ex)
    if (aligned) { if (cj) GO(true, true); else GO(false, true); }
    else         { if (cj) GO(true, false); else GO(false, false); }
#undef GO
    return check_launch("dot_mg_kernel");
}

// ----------------------------------------------------------------------------------------------------------------
// accblas_dot_static: the dot for callers that must not allocate
// ----------------------------------------------------------------------------------------------------------------
// The per-CTA partial sums of the two-kernel dot live in an array that is part of the library image instead of in
// memory from a pool: loading libaccblas_b200.so on a GPU reserves it (ACCBLAS_DOT_STATIC_SLOTS x 32 bytes = 32 KiB)
// and it stays until the process exits.  g_dot_static_scratch is a device symbol: device code may name it, host
// code may NOT use the name as a pointer -- the host sees a stand-in object, and a kernel handed that would read
// host memory or fault.  dot_static_scratch() asks the runtime for the real address, once per device.
constexpr int DOT_STATIC_SLOT_BYTES = 32;
static __device__ __align__(256) unsigned char g_dot_static_scratch[ACCBLAS_DOT_STATIC_SLOTS * DOT_STATIC_SLOT_BYTES];

// Device address of the array on the CURRENT device (each device that loads the module has its own copy).  The look-up
// is the one piece of set-up this entry point has; it makes the runtime load the module on that device if it has not
// yet, which is why a first call per device belongs before any stream capture.  Process-wide table, not per thread:
// the address belongs to the device, whichever host thread asks.
int dot_static_scratch(void **out)
{
    static std::atomic<void *> addr[64];
    int dev = 0;
    ACC_CUDA(cudaGetDevice(&dev));
    if (dev < 0 || dev >= 64) return cuda_fail(cudaErrorInvalidDevice, "accblas_dot_static: device ordinal beyond 63");
    void *p = addr[dev].load(std::memory_order_acquire);
    if (!p) {
        ACC_CUDA(cudaGetSymbolAddress(&p, g_dot_static_scratch));
        if (!p) return cuda_fail(cudaErrorInvalidSymbol, "cudaGetSymbolAddress(g_dot_static_scratch)");
        addr[dev].store(p, std::memory_order_release);      // two threads racing here store the same value
    }
    *out = p;
    return ACCBLAS_OK;
}

// Same CTA shape, grid and summation tree as accblas_dot's default path (the last CTA of dot_mg_kernel and
// dot_final_kernel<.., THREADS> add the partials in the same order), so the two entry points return the same bits.
template <typename T, typename P, int THREADS, int UNROLL, int MINB>
int dot_static_launch(cudaStream_t st, bool conj, int64_t n, const T *x, const T *y, T *out)
{
    using Part = typename P::Part;
    static_assert(sizeof(Part) <= DOT_STATIC_SLOT_BYTES, "a partial fits its slot");
    const bool aligned = both_aligned(x, y);
    const int grid = grid_for<T>(n, aligned, THREADS, MINB);
    if (grid > ACCBLAS_DOT_STATIC_SLOTS)
        return fail_text(ACCBLAS_ERR_WORKSPACE,
                         "accblas_dot_static: this device wants more CTAs than the built-in scratch has slots "
                         "(ACCBLAS_DOT_STATIC_SLOTS); use accblas_dot");
    void *scratch = nullptr;
    ACC_TRY(dot_static_scratch(&scratch));
    Part *partial = static_cast<Part *>(scratch);
    const bool cj = conj && Scalar<T>::cplx;
    // every slot the second kernel reads is written by the first: nothing in the array has to start at zero
#define GO(CJ, AL) dot_partial_kernel<T, P, CJ, AL, THREADS, UNROLL, MINB><<<grid, THREADS, 0, st>>>(n, x, y, partial)
    if (aligned) { if (cj) GO(true, true); else GO(false, true); }
    else         { if (cj) GO(true, false); else GO(false, false); }
#undef GO
    ACC_TRY(check_launch("dot_partial_kernel (static scratch)"));
    dot_final_kernel<T, P, THREADS><<<1, THREADS, 0, st>>>(grid, partial, out);
    return check_launch("dot_final_kernel (static scratch)");
}

template <typename T>
int dot_static_typed(cudaStream_t st, bool conj, int64_t n, const T *x, const T *y, T *out)
{
    if (n == 0) {   // fold (+) 0 over nothing
        ACC_CUDA(cudaMemsetAsync(out, 0, sizeof(T), st));
        return ACCBLAS_OK;
    }
    if (sizeof(T) == 4) return dot_static_launch<T, Plain<T, 8>, 256, 8, 2>(st, conj, n, x, y, out);
    return dot_static_launch<T, Plain<T, 4>, 512, 4, 2>(st, conj, n, x, y, out);
}

}  // namespace

int dot_static_dispatch(cudaStream_t s, int dtype, int conj_x, int64_t n, const void *x, const void *y, void *out)
{
    switch (dtype) {
    case ACCBLAS_F32: return dot_static_typed<float>(s, false, n, (const float *)x, (const float *)y, (float *)out);
    case ACCBLAS_F64: return dot_static_typed<double>(s, false, n, (const double *)x, (const double *)y, (double *)out);
    case ACCBLAS_C32: return dot_static_typed<float2>(s, conj_x != 0, n, (const float2 *)x, (const float2 *)y, (float2 *)out);
    default:          return dot_static_typed<double2>(s, conj_x != 0, n, (const double2 *)x, (const double2 *)y, (double2 *)out);
    }
}

int dot_dispatch(cudaStream_t s, int dtype, int conj_x, int64_t n, const void *x, const void *y, void *out, bool compensated)
{
    switch (dtype) {
    case ACCBLAS_F32: return dot_typed<float>(s, false, n, (const float *)x, (const float *)y, (float *)out, compensated);
    case ACCBLAS_F64: return dot_typed<double>(s, false, n, (const double *)x, (const double *)y, (double *)out, compensated);
    case ACCBLAS_C32: return dot_typed<float2>(s, conj_x != 0, n, (const float2 *)x, (const float2 *)y, (float2 *)out, compensated);
    default:          return dot_typed<double2>(s, conj_x != 0, n, (const double2 *)x, (const double2 *)y, (double2 *)out, compensated);
    }
}

int dot_mg_dispatch(cudaStream_t s, int dtype, int conj_x, int64_t n, const void *x, const void *y, void *out,
                    const DotExchange &ex, int max_grid)
{
    switch (dtype) {
    case ACCBLAS_F32: return dot_mg_typed<float>(s, false, n, (const float *)x, (const float *)y, (float *)out, ex, max_grid);
    case ACCBLAS_F64: return dot_mg_typed<double>(s, false, n, (const double *)x, (const double *)y, (double *)out, ex, max_grid);
    case ACCBLAS_C32: return dot_mg_typed<float2>(s, conj_x != 0, n, (const float2 *)x, (const float2 *)y, (float2 *)out, ex, max_grid);
    default:          return dot_mg_typed<double2>(s, conj_x != 0, n, (const double2 *)x, (const double2 *)y, (double2 *)out, ex, max_grid);
    }
}

}  // namespace accblas

// See include/accblas_b200.h.  Argument checks come first and touch no CUDA state, so the limits can be probed on a
// machine without a GPU.
extern "C" int accblas_dot_static(accblas_stream_t stream, int dtype, int conj_x, int64_t n, const void *x, const void *y,
                                  void *out_dev)
{
    using namespace accblas;
    if (dtype != ACCBLAS_F32 && dtype != ACCBLAS_F64 && dtype != ACCBLAS_C32 && dtype != ACCBLAS_C64) return ACCBLAS_ERR_DTYPE;
    if (n < 0) return ACCBLAS_ERR_SHAPE;
    if (n > ACCBLAS_DOT_STATIC_MAX_N)
        return fail_text(ACCBLAS_ERR_WORKSPACE,
                         "accblas_dot_static: n is beyond ACCBLAS_DOT_STATIC_MAX_N (2^32 elements); use accblas_dot");
    if (!out_dev) return ACCBLAS_ERR_NULL;
    if (n > 0 && (!x || !y)) return ACCBLAS_ERR_NULL;
    ACC_TRY(device_check());
    return dot_static_dispatch(static_cast<cudaStream_t>(stream), dtype, conj_x, n, x, y, out_dev);
}
