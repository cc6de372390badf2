// common.cuh -- shared device/host helpers for libaccblas_b200.so (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stddef.h>
#include <atomic>
#include "../../include/accblas_b200.h"

namespace accblas {

// ----------------------------------------------------------------------------------
// per-thread bookkeeping (the library itself is stateless: these are diagnostics)
// ----------------------------------------------------------------------------------
extern thread_local uint64_t g_launches;      // kernels enqueued by this thread
extern std::atomic<uint64_t> g_launches_total; // ... by every thread of the process
extern thread_local char g_cuda_err[256];     // text of the last CUDA failure
extern thread_local int g_gemm_path;          // accblas_set_gemm_path
// accblas_set_tuning (-1 = automatic).  Per thread; a new thread starts from the process-wide defaults
// in the environment variable ACCBLAS_TUNE ("knob=value,knob=value", e.g. "7=0,9=0" pins 4M complex
// products) -- the way to pin a choice from a host runtime whose FFI calls hop between OS threads.
struct Tuning {
    int v[ACCBLAS_TUNE_COUNT];
    Tuning();
    int &operator[](int i) { return v[i]; }
};
extern thread_local Tuning g_tuning;

int cuda_fail(cudaError_t e, const char *where);   // records text, returns ACCBLAS_ERR_CUDA
int fail_text(int status, const char *text);       // records text for a non-CUDA failure, returns status
inline int check_launch(const char *where) {
    ++g_launches;
    g_launches_total.fetch_add(1, std::memory_order_relaxed);
    cudaError_t e = cudaPeekAtLastError();
    return e == cudaSuccess ? ACCBLAS_OK : cuda_fail(cudaGetLastError(), where);
}
#define ACC_CUDA(call)                                                     \
    do {                                                                   \
        cudaError_t _e = (call);                                           \
        if (_e != cudaSuccess) return ::accblas::cuda_fail(_e, #call);     \
    } while (0)
#define ACC_TRY(call)                                                      \
    do {                                                                   \
        int _s = (call);                                                   \
        if (_s != ACCBLAS_OK) return _s;                                   \
    } while (0)

// cudaFuncSetAttribute (dynamic shared memory opt-in) is per DEVICE: a `static thread_local AttrOnce` beside each
// kernel instantiation remembers on which devices this thread has already done it, so a thread that serves
// several GPUs sets the attribute once on each.
struct AttrOnce {
    uint64_t mask = 0;
    bool need() {
        int d = 0;
        if (cudaGetDevice(&d) != cudaSuccess || d < 0 || d > 63) return true;
        if ((mask >> d) & 1u) return false;
        mask |= 1ull << d;
        return true;
    }
};

int sm_count();                                  // SMs of the current device (cached per device)
// stream-ordered scratch from the library's OWN memory pool (api.cu); release with cudaFreeAsync(p, stream)
cudaError_t scratch_alloc(void **p, size_t bytes, cudaStream_t stream);
int device_check();                              // ACCBLAS_OK iff current device is CC 10.x

inline size_t elem_size(int dtype) {
    return dtype == ACCBLAS_F32 ? 4 : dtype == ACCBLAS_C64 ? 16 : 8;
}
inline bool is_complex(int dtype) { return dtype == ACCBLAS_C32 || dtype == ACCBLAS_C64; }

// ----------------------------------------------------------------------------------
// scalar algebra over the four element types: float, double, float2, double2
// ----------------------------------------------------------------------------------
template <typename T> struct Scalar;
template <> struct Scalar<float> {
    using real = float; static constexpr bool cplx = false;
    __host__ __device__ static inline float zero() { return 0.f; }
};
template <> struct Scalar<double> {
    using real = double; static constexpr bool cplx = false;
    __host__ __device__ static inline double zero() { return 0.0; }
};
template <> struct Scalar<float2> {
    using real = float; static constexpr bool cplx = true;
    __host__ __device__ static inline float2 zero() { return make_float2(0.f, 0.f); }
};
template <> struct Scalar<double2> {
    using real = double; static constexpr bool cplx = true;
    __host__ __device__ static inline double2 zero() { return make_double2(0.0, 0.0); }
};

__device__ __forceinline__ float  fma_acc(float a, float b, float c) { return fmaf(a, b, c); }
__device__ __forceinline__ double fma_acc(double a, double b, double c) { return fma(a, b, c); }
__device__ __forceinline__ float2 fma_acc(float2 a, float2 b, float2 c) {
    c.x = fmaf(a.x, b.x, c.x); c.x = fmaf(-a.y, b.y, c.x);
    c.y = fmaf(a.x, b.y, c.y); c.y = fmaf(a.y, b.x, c.y);
    return c;
}
__device__ __forceinline__ double2 fma_acc(double2 a, double2 b, double2 c) {
    c.x = fma(a.x, b.x, c.x); c.x = fma(-a.y, b.y, c.x);
    c.y = fma(a.x, b.y, c.y); c.y = fma(a.y, b.x, c.y);
    return c;
}
__device__ __forceinline__ float  mul_s(float a, float b) { return a * b; }
__device__ __forceinline__ double mul_s(double a, double b) { return a * b; }
__device__ __forceinline__ float2 mul_s(float2 a, float2 b) {
    return make_float2(a.x * b.x - a.y * b.y, a.x * b.y + a.y * b.x);
}
__device__ __forceinline__ double2 mul_s(double2 a, double2 b) {
    return make_double2(a.x * b.x - a.y * b.y, a.x * b.y + a.y * b.x);
}
__device__ __forceinline__ float  add_s(float a, float b) { return a + b; }
__device__ __forceinline__ double add_s(double a, double b) { return a + b; }
__device__ __forceinline__ float2 add_s(float2 a, float2 b) { return make_float2(a.x + b.x, a.y + b.y); }
__device__ __forceinline__ double2 add_s(double2 a, double2 b) { return make_double2(a.x + b.x, a.y + b.y); }
__device__ __forceinline__ float  conj_s(float a) { return a; }
__device__ __forceinline__ double conj_s(double a) { return a; }
__device__ __forceinline__ float2 conj_s(float2 a) { return make_float2(a.x, -a.y); }
__device__ __forceinline__ double2 conj_s(double2 a) { return make_double2(a.x, -a.y); }

__device__ __forceinline__ float  shfl_xor_s(float v, int m) { return __shfl_xor_sync(0xffffffffu, v, m); }
__device__ __forceinline__ double shfl_xor_s(double v, int m) { return __shfl_xor_sync(0xffffffffu, v, m); }
__device__ __forceinline__ float2 shfl_xor_s(float2 v, int m) {
    return make_float2(__shfl_xor_sync(0xffffffffu, v.x, m), __shfl_xor_sync(0xffffffffu, v.y, m));
}
__device__ __forceinline__ double2 shfl_xor_s(double2 v, int m) {
    return make_double2(__shfl_xor_sync(0xffffffffu, v.x, m), __shfl_xor_sync(0xffffffffu, v.y, m));
}
template <typename T> __device__ __forceinline__ T warp_sum(T v) {
#pragma unroll
    for (int m = 16; m > 0; m >>= 1) v = add_s(v, shfl_xor_s(v, m));
    return v;
}

// 128-bit streaming load (read once: do not allocate in L1)
__device__ __forceinline__ int4 ldg_stream(const void *p) {
    int4 r;
    asm volatile("ld.global.nc.L1::no_allocate.v4.s32 {%0,%1,%2,%3}, [%4];"
                 : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p));
    return r;
}
// 128-bit cached load (re-used operand such as the gemv vector)
__device__ __forceinline__ int4 ldg_cached(const void *p) {
    return __ldg(reinterpret_cast<const int4 *>(p));
}

// alpha as a kernel argument: a value read on the host at call time (the reference's convention,
// LLVM/PTX/Level3.hs:81,89) or a device pointer dereferenced by the kernel (no host round trip).
template <typename T> struct Alpha {
    T v;
    const T *dev;
    __device__ __forceinline__ T get() const { return dev ? *dev : v; }
};
template <typename T> inline Alpha<T> make_alpha(const void *p, bool on_device) {
    Alpha<T> a;
    if (on_device) { a.v = Scalar<T>::zero(); a.dev = static_cast<const T *>(p); }
    else { a.v = *static_cast<const T *>(p); a.dev = nullptr; }
    return a;
}

// beta * Cin term of the extended product (accblas_gemm_ex): D = alpha * op(A) * op(B) + beta * Cin.
// cin == nullptr means beta = 0: Cin is never read (the reference's only mode, LLVM/PTX/Level3.hs:90).
// Cin may be the output itself (in place: every element is read and then written by the same thread) or a
// separate array with its own row pitch (e.g. a local accumulator while the output lives in a peer's HBM).
template <typename T> struct BetaC {
    const T *cin;
    int64_t ld;
    T beta;
    __device__ __forceinline__ bool on() const { return cin != nullptr; }
    // v + beta * Cin[r][c]
    __device__ __forceinline__ T add(T v, int64_t r, int64_t c) const { return fma_acc(beta, cin[r * ld + c], v); }
};

// ----------------------------------------------------------------------------------
// kernel-family entry points (one translation unit each)
// ----------------------------------------------------------------------------------
struct GemmArgs {
    int dtype, opA, opB;
    int64_t rowsA, colsA, rowsB, colsB;   // stored (pre-op) shapes
    int64_t m, n, k;                      // derived: C is m x n, inner extent k
    int64_t lda, ldb, ldc;                // row pitches in ELEMENTS (0 = tight: colsA, colsB, n)
    const void *alpha;                    // one element: host pointer, or device pointer if alpha_dev
    bool alpha_dev;                       // alpha is read by the kernel (accblas_*_alpha_dev entry points)
    const void *A, *B;
    void *C;
    cudaStream_t stream;
    // accblas_gemm_ex: D = alpha * op(A) * op(B) + beta * Cin.  beta == nullptr (or *beta == 0): Cin is never read.
    const void *beta;                     // one element in host memory
    const void *Cin;                      // m x n, row pitch ldcin (elements; 0 = ldc); may equal C
    int64_t ldcin;
    bool no_persistent;                   // never the persistent (one CTA per SM) schedule: set by callers that run several
                                          // products CONCURRENTLY on different streams (the host wavefront)
    // strided batch (accblas_gemm_strided_batched): `batch` independent products, operand b of the batch at
    // base + b * stride (elements; stride 0 = the same operand for every product).  batch <= 1: one product.
    int64_t batch, strideA, strideB, strideC;
    // internal (gemm_dmma.cu): split-K over idle SMs -- ksplit slices of the K-blocks, slice s writes its alpha-scaled partial
    // product to ksplit_ws + s * m * n (row pitch n); a reduce kernel sums the slices into C.  0 / 1: no split.
    int ksplit;
    void *ksplit_ws;
};

template <typename T> inline bool scalar_is_zero(const T &v);
template <> inline bool scalar_is_zero<float>(const float &v) { return v == 0.f; }
template <> inline bool scalar_is_zero<double>(const double &v) { return v == 0.0; }
template <> inline bool scalar_is_zero<float2>(const float2 &v) { return v.x == 0.f && v.y == 0.f; }
template <> inline bool scalar_is_zero<double2>(const double2 &v) { return v.x == 0.0 && v.y == 0.0; }
template <typename T> inline BetaC<T> make_beta(const GemmArgs &g) {
    BetaC<T> b;
    b.cin = nullptr; b.ld = 0; b.beta = Scalar<T>::zero();
    if (g.beta && g.Cin) {
        const T v = *static_cast<const T *>(g.beta);
        if (!scalar_is_zero<T>(v)) { b.cin = static_cast<const T *>(g.Cin); b.ld = g.ldcin ? g.ldcin : g.ldc; b.beta = v; }
    }
    return b;
}
inline bool has_beta(const GemmArgs &g) {
    if (!g.beta || !g.Cin) return false;
    switch (g.dtype) {
    case ACCBLAS_F32: return !scalar_is_zero(*static_cast<const float *>(g.beta));
    case ACCBLAS_F64: return !scalar_is_zero(*static_cast<const double *>(g.beta));
    case ACCBLAS_C32: return !scalar_is_zero(*static_cast<const float2 *>(g.beta));
    default:          return !scalar_is_zero(*static_cast<const double2 *>(g.beta));
    }
}

int gemm_device(GemmArgs &g);                         // api.cu : validate + dispatch one device-resident product
int gemm_host_pipeline(const GemmArgs &g, const void *A_host, const void *B_host,
                       void *C_host);                 // host_pipeline.cu : upload / multiply / download wavefront
int gemm_generic(const GemmArgs &g);                  // gemm_generic.cu : any shape
int gemm_rank1(const GemmArgs &g);
int splitk_reduce(const GemmArgs &g);                //                   sum of the split-K slices (ksplit_ws -> C)                    // gemm_generic.cu : k == 1 (outer product `><`)
bool dgemm_dmma_eligible(const GemmArgs &g);          // gemm_dmma.cu    : f64 / c64 on DMMA
int  dgemm_dmma(const GemmArgs &g);
bool dgemm_direct_eligible(const GemmArgs &g);        //                   f64, misaligned operands, one wave of 32 x 32 tiles, whole K in smem
int  dgemm_direct(const GemmArgs &g);
size_t dgemm_dmma_workspace(const GemmArgs &g);        //                   stream-K scratch (0 for the data-parallel tiles)
bool zgemm_3m_eligible(const GemmArgs &g);            // gemm_3m.cu      : c64 as three real DMMA products
bool zgemm_3m_wanted(const GemmArgs &g);              //                   (knob + cost model)
int  zgemm_3m(const GemmArgs &g);
size_t zgemm_3m_workspace(const GemmArgs &g);
bool sgemm_tc_eligible(const GemmArgs &g);            // gemm_tf32x3.cu  : f32 / c32 on tcgen05
int  sgemm_tc(const GemmArgs &g);
size_t sgemm_tc_workspace(const GemmArgs &g);
bool cgemm_3m_wanted(const GemmArgs &g);              //                   c32 as three real tcgen05 products
int  cgemm_3m(const GemmArgs &g);
size_t cgemm_3m_workspace(const GemmArgs &g);
bool sgemm_mma_eligible(const GemmArgs &g);           // gemm_mma_tf32.cu : f32 / c32 on mma.sync, small / medium
int  sgemm_mma(const GemmArgs &g);
bool sgemm_mma_direct_eligible(const GemmArgs &g);    //                    f32, misaligned operands, whole K in shared memory
int  sgemm_mma_direct(const GemmArgs &g);

int gemv_dispatch(cudaStream_t s, int dtype, int opA, int64_t rowsA, int64_t colsA,
                  const void *alpha, bool alpha_dev, const void *A, const void *x, void *y);   // gemv.cu
int dot_dispatch(cudaStream_t s, int dtype, int conj_x, int64_t n,
                 const void *x, const void *y, void *out, bool compensated = false);  // dot.cu

int sdot_dispatch(cudaStream_t s, int dtype, int64_t n, const void *z_host,
                  const void *x, const void *y, void *out);                           // level1.cu
int asum_dispatch(cudaStream_t s, int dtype, int64_t n, const void *x, void *out);
int iamax_dispatch(cudaStream_t s, int dtype, int want_max, int64_t n, const void *x, int64_t *out);

}  // namespace accblas
