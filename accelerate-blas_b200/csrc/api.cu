// api.cu -- extern "C" surface of libaccblas_b200.so (include/accblas_b200.h):
// argument validation, shape derivation, kernel-family dispatch, host-buffer entry points.
// There is NO CPU fallback anywhere in this file: every path ends in a kernel launch on
// the caller's stream or in an error status.
#include "common.cuh"
#include <vector>
#include <cstdio>
#include <cstring>
#include <cstdlib>
#include <algorithm>
#include <atomic>
#include <mutex>

namespace accblas {

std::atomic<uint64_t> g_launches_total{0};    // every thread's launches (accblas_launch_count_total)
thread_local uint64_t g_launches = 0;
thread_local char g_cuda_err[256] = {0};
thread_local int g_gemm_path = 0;
namespace {
struct TuningDefaults {
    int v[ACCBLAS_TUNE_COUNT];
    TuningDefaults()
    {
        for (int &x : v) x = -1;
        const char *e = getenv("ACCBLAS_TUNE");
        while (e && *e) {                                   // "knob=value,knob=value"
            char *end = nullptr;
            const long knob = strtol(e, &end, 10);
            if (end == e || *end != '=') break;
            e = end + 1;
            const long val = strtol(e, &end, 10);
            if (end == e) break;
            if (knob >= 0 && knob < ACCBLAS_TUNE_COUNT) v[knob] = (int)val;
            e = *end == ',' ? end + 1 : end;
        }
    }
};
}  // namespace
Tuning::Tuning()
{
    static const TuningDefaults d;
    for (int i = 0; i < ACCBLAS_TUNE_COUNT; ++i) v[i] = d.v[i];
}
thread_local Tuning g_tuning;

namespace {
std::mutex g_err_mu;
char g_cuda_err_any[256] = {0};               // most recent failure of ANY thread (accblas_last_cuda_error_r)
}  // namespace

int cuda_fail(cudaError_t e, const char *where)
{
    snprintf(g_cuda_err, sizeof(g_cuda_err), "%s: %s (%s)", where, cudaGetErrorName(e), cudaGetErrorString(e));
    {
        std::lock_guard<std::mutex> lock(g_err_mu);
        memcpy(g_cuda_err_any, g_cuda_err, sizeof(g_cuda_err_any));
    }
    return ACCBLAS_ERR_CUDA;
}

// A failure that is not a CUDA error but still deserves a sentence (a built-in limit exceeded): the text goes where
// accblas_last_cuda_error / accblas_last_cuda_error_r find it, the status is returned unchanged.
int fail_text(int status, const char *text)
{
    snprintf(g_cuda_err, sizeof(g_cuda_err), "%s", text);
    {
        std::lock_guard<std::mutex> lock(g_err_mu);
        memcpy(g_cuda_err_any, g_cuda_err, sizeof(g_cuda_err_any));
    }
    return status;
}

namespace {
struct DevInfo { int sms = 0; int cc_major = 0; };
DevInfo &dev_info()
{
    static thread_local DevInfo cache[64];
    static thread_local DevInfo none;
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return none;
    DevInfo &d = cache[dev];
    if (d.sms == 0) {
        cudaDeviceGetAttribute(&d.sms, cudaDevAttrMultiProcessorCount, dev);
        cudaDeviceGetAttribute(&d.cc_major, cudaDevAttrComputeCapabilityMajor, dev);
    }
    return d;
}

// Stream-ordered scratch comes from a memory pool the LIBRARY owns (one per device, created on first use), not
// from the device's default pool: inside an Accelerate process the default pool belongs to everyone who calls
// cudaMallocAsync, and the library must not change its release threshold.  The private pool keeps freed blocks
// cached (release threshold = max: otherwise every call pays a physical allocation, ~0.2 ms) until
// accblas_trim_workspace() hands them back to the driver.
std::mutex g_pool_mu;
cudaMemPool_t g_pools[64] = {nullptr};

cudaError_t scratch_pool(cudaMemPool_t *out)
{
    int dev = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return e;
    if (dev < 0 || dev >= 64) return cudaErrorInvalidDevice;
    std::lock_guard<std::mutex> lock(g_pool_mu);
    if (!g_pools[dev]) {
        cudaMemPoolProps props{};
        props.allocType = cudaMemAllocationTypePinned;
        props.handleTypes = cudaMemHandleTypeNone;
        props.location.type = cudaMemLocationTypeDevice;
        props.location.id = dev;
        cudaMemPool_t pool = nullptr;
        e = cudaMemPoolCreate(&pool, &props);
        if (e != cudaSuccess) return e;
        uint64_t keep = ~0ull;
        cudaMemPoolSetAttribute(pool, cudaMemPoolAttrReleaseThreshold, &keep);
        g_pools[dev] = pool;
    }
    *out = g_pools[dev];
    return cudaSuccess;
}
}  // namespace

cudaError_t scratch_alloc(void **p, size_t bytes, cudaStream_t s)
{
    static const bool use_default = [] { const char *e = getenv("ACCBLAS_DEFAULT_POOL"); return e && *e == '1'; }();
    if (use_default) return cudaMallocAsync(p, bytes ? bytes : 1, s);     // debugging aid: the device's default pool
    cudaMemPool_t pool = nullptr;
    cudaError_t e = scratch_pool(&pool);
    if (e != cudaSuccess) return e;
    return cudaMallocFromPoolAsync(p, bytes ? bytes : 1, pool, s);
}

int sm_count() { int s = dev_info().sms; return s > 0 ? s : 148; }
int device_check() { return dev_info().cc_major == 10 ? ACCBLAS_OK : ACCBLAS_ERR_ARCH; }

namespace {

bool bad_dtype(int d) { return d < ACCBLAS_F32 || d > ACCBLAS_C64; }
bool has_beta_value(int dtype, const void *beta)
{
    GemmArgs t{};
    t.dtype = dtype; t.beta = beta; t.Cin = beta;      // any non-null Cin: only the VALUE is looked at
    return has_beta(t);
}
bool bad_op(int o) { return o < ACCBLAS_OP_N || o > ACCBLAS_OP_H; }

bool overlaps(const void *a, size_t na, const void *b, size_t nb)
{
    if (!a || !b || na == 0 || nb == 0) return false;
    const uintptr_t a0 = (uintptr_t)a, b0 = (uintptr_t)b;
    return a0 < b0 + nb && b0 < a0 + na;
}

// elements spanned by a rows x cols view with row pitch ld
size_t span(int64_t rows, int64_t cols, int64_t ld)
{
    return rows > 0 && cols > 0 ? (size_t)(rows - 1) * (size_t)ld + (size_t)cols : 0;
}

// shape derivation shared by the device and host entry points (LLVM/PTX/Level3.hs:54-65)
int derive_gemm(GemmArgs &g)
{
    if (bad_dtype(g.dtype)) return ACCBLAS_ERR_DTYPE;
    if (bad_op(g.opA) || bad_op(g.opB)) return ACCBLAS_ERR_OP;
    if (g.rowsA < 0 || g.colsA < 0 || g.rowsB < 0 || g.colsB < 0) return ACCBLAS_ERR_SHAPE;
    g.m = g.opA == ACCBLAS_OP_N ? g.rowsA : g.colsA;
    g.k = g.opA == ACCBLAS_OP_N ? g.colsA : g.rowsA;
    g.n = g.opB == ACCBLAS_OP_N ? g.colsB : g.rowsB;
    const int64_t kb = g.opB == ACCBLAS_OP_N ? g.rowsB : g.colsB;
    if (g.k != kb) return ACCBLAS_ERR_SHAPE;     // the reference never checks this
    // row pitches: 0 = tight (what the reference always passes, LLVM/PTX/Level3.hs:64-65,91)
    if (g.lda == 0) g.lda = g.colsA;
    if (g.ldb == 0) g.ldb = g.colsB;
    if (g.ldc == 0) g.ldc = g.n;
    if (g.lda < g.colsA || g.ldb < g.colsB || g.ldc < g.n) return ACCBLAS_ERR_SHAPE;
    if (!is_complex(g.dtype)) {                  // H on a real type is T (BLAS/Level3.hs:100,110)
        if (g.opA == ACCBLAS_OP_H) g.opA = ACCBLAS_OP_T;
        if (g.opB == ACCBLAS_OP_H) g.opB = ACCBLAS_OP_T;
    }
    return ACCBLAS_OK;
}

enum Family { FAM_NONE, FAM_GENERIC, FAM_DMMA, FAM_TC, FAM_RANK1, FAM_MMA, FAM_Z3M, FAM_C3M };

// f32 / c32: the tcgen05 path (pack pass + 128 x 256 tiles) wins once the product fills the GPU with
// such tiles; below that the one-launch mma.sync path does (profiles/r01_sgemm_families.jsonl)
constexpr double kTcMinWork = 7.0e8;     // real-equivalent m * n * k (between 768^3 and 1000^3 since the 128 x 64 tile)

Family pick_family(const GemmArgs &g)
{
    if (g.m == 0 || g.n == 0) return FAM_NONE;
    if (g_gemm_path == 1) return FAM_GENERIC;
    if (g.k == 1 && g_gemm_path == 0) return FAM_RANK1;          // `><`: a store stream, not a contraction
    if ((g.dtype == ACCBLAS_F64 || g.dtype == ACCBLAS_C64) && dgemm_dmma_eligible(g))
        return g.dtype == ACCBLAS_C64 && zgemm_3m_wanted(g) ? FAM_Z3M : FAM_DMMA;
    if (g.dtype == ACCBLAS_F32 || g.dtype == ACCBLAS_C32) {
        const bool tc = sgemm_tc_eligible(g), mma = sgemm_mma_eligible(g);
        const int want = g_tuning[ACCBLAS_TUNE_SGEMM_FAMILY];
        if (want == 1 && mma) return FAM_MMA;
        if (want == 2 && tc) return cgemm_3m_wanted(g) ? FAM_C3M : FAM_TC;
        const double work = (g.dtype == ACCBLAS_C32 ? 4.0 : 1.0) * (double)g.m * (double)g.n * (double)g.k;
        if (tc && !mma && work < kTcMinWork && want != 2 && g_gemm_path == 0) {
            // Small product whose operands only miss TMA's alignment (250 x 250 x 512 Floats): the pack pass + tcgen05 tiles
            // measured 24.6 us there against cuBLAS 10.3; the mma.sync family serves it -- through its direct variant or behind
            // the re-pitch copy (gemm_repitched) -- if it would take the aligned product.
            GemmArgs h = g;
            const int64_t q = g.dtype == ACCBLAS_C32 ? 2 : 4;
            h.A = h.B = nullptr;
            h.lda = (g.lda + q - 1) / q * q; h.ldb = (g.ldb + q - 1) / q * q;
            if (sgemm_mma_eligible(h)) return FAM_GENERIC;
        }
        if (tc && (work >= kTcMinWork || !mma)) return cgemm_3m_wanted(g) ? FAM_C3M : FAM_TC;
        if (mma) return FAM_MMA;
    }
    return FAM_GENERIC;
}

int run_family(Family f, const GemmArgs &g);
int gemm_repitched(const GemmArgs &g);
}  // namespace

int gemm_device(GemmArgs &g)
{
    ACC_TRY(derive_gemm(g));
    if (g.m == 0 || g.n == 0) return ACCBLAS_OK;                       // nothing to write
    if (!g.C || !g.alpha) return ACCBLAS_ERR_NULL;
    const size_t es = elem_size(g.dtype);
    const size_t bytesC = es * span(g.m, g.n, g.ldc);
    if (has_beta(g)) {                                                 // checked first: the k == 0 product reads Cin too
        // Cin is either the output itself (same base AND pitch: in place) or an array that does not touch it
        if (g.ldcin == 0) g.ldcin = g.ldc;
        if (g.ldcin < g.n) return ACCBLAS_ERR_SHAPE;
        const bool in_place = g.Cin == g.C && g.ldcin == g.ldc;
        if (!in_place && overlaps(g.C, bytesC, g.Cin, es * span(g.m, g.n, g.ldcin))) return ACCBLAS_ERR_ALIAS;
    } else {
        g.beta = nullptr; g.Cin = nullptr;
    }
    if (g.k == 0 && !has_beta(g)) {                                    // alpha * (empty sum) = 0
        ACC_CUDA(cudaMemset2DAsync(g.C, es * (size_t)g.ldc, 0, es * (size_t)g.n, (size_t)g.m, g.stream));
        return ACCBLAS_OK;
    }
    if (g.k == 0) {                                                    // D = beta * Cin: the SIMT kernel's epilogue alone
        ACC_TRY(device_check());
        return gemm_generic(g);
    }
    if (!g.A || !g.B) return ACCBLAS_ERR_NULL;
    if (overlaps(g.C, bytesC, g.A, es * span(g.rowsA, g.colsA, g.lda)) ||
        overlaps(g.C, bytesC, g.B, es * span(g.rowsB, g.colsB, g.ldb)))
        return ACCBLAS_ERR_ALIAS;
    ACC_TRY(device_check());
    const Family f = pick_family(g);
    if (g_gemm_path == 2 && f == FAM_GENERIC) return ACCBLAS_ERR_SHAPE;
    if (f == FAM_GENERIC && g_gemm_path == 0 && (double)g.m * (double)g.n * (double)g.k >= 48.0 * 48.0 * 48.0)
        return gemm_repitched(g);
    return run_family(f, g);
}

namespace {
int run_family(Family f, const GemmArgs &g)
{
    switch (f) {
    case FAM_DMMA: return dgemm_dmma(g);
    case FAM_TC:   return sgemm_tc(g);
    case FAM_RANK1: return gemm_rank1(g);
    case FAM_MMA:  return sgemm_mma(g);
    case FAM_Z3M:  return zgemm_3m(g);
    case FAM_C3M:  return cgemm_3m(g);
    default:       return gemm_generic(g);
    }
}

struct RepitchArgs {
    const char *src; char *dst;
    int64_t rows, words, src_pitch, dst_pitch;      // extents and pitches in copy words (rows == 0: nothing to do)
};
template <typename W>
__global__ void __launch_bounds__(256) repitch_kernel(const RepitchArgs a, const RepitchArgs b)
{
    const RepitchArgs &r = blockIdx.z == 0 ? a : b;
    const int64_t c = (int64_t)blockIdx.x * 256 + threadIdx.x;
    if (c >= r.words) return;
    const W *src = reinterpret_cast<const W *>(r.src);
    W *dst = reinterpret_cast<W *>(r.dst);
    for (int64_t row = blockIdx.y; row < r.rows; row += gridDim.y) dst[row * r.dst_pitch + c] = src[row * r.src_pitch + c];
}

// The TMA-fed families need 16-byte aligned bases and row pitches.  Dense Accelerate arrays with an
// odd column count (the reference benchmarks 250 x 250 Floats: 1000-byte rows) miss that by a few
// bytes, which used to send them to the SIMT kernel at half cuBLAS's speed.  Instead: copy the
// offending operand(s) once into stream-ordered scratch with the pitch rounded up (an O(n^2) strided
// D2D copy on the caller's stream) and run the tensor family on the copy.  C keeps its own layout
// (the epilogues fall back to scalar stores); a misaligned C base stays on the SIMT kernel.
int gemm_repitched(const GemmArgs &g)
{
    const size_t es = elem_size(g.dtype);
    const int64_t q = (int64_t)(16 / es);                              // elements per 16 bytes
    auto bad = [&](const void *p, int64_t ld) { return (reinterpret_cast<uintptr_t>(p) & 15u) != 0 || (ld % q) != 0; };
    const bool fixA = bad(g.A, g.lda), fixB = bad(g.B, g.ldb);
    if ((!fixA && !fixB) || (reinterpret_cast<uintptr_t>(g.C) & 15u) != 0) return gemm_generic(g);
    // small f32 products: the mma.sync kernel loads the misaligned panels itself (one launch, no copy, no scratch)
    if (g_tuning[ACCBLAS_TUNE_SGEMM_DIRECT] != 0 && sgemm_mma_direct_eligible(g)) return sgemm_mma_direct(g);
    if (g.dtype == ACCBLAS_F64 && dgemm_direct_eligible(g)) return dgemm_direct(g);
    GemmArgs h = g;
    const int64_t lda = fixA ? (g.colsA + q - 1) / q * q : g.lda, ldb = fixB ? (g.colsB + q - 1) / q * q : g.ldb;
    h.lda = lda; h.ldb = ldb;
    if (fixA) h.A = nullptr;                                            // "aligned" for the eligibility test
    if (fixB) h.B = nullptr;
    const Family f = pick_family(h);
    if (f == FAM_GENERIC || f == FAM_NONE) return gemm_generic(g);
    const size_t bytesA = fixA ? (es * (size_t)lda * (size_t)g.rowsA + 255) / 256 * 256 : 0;
    const size_t bytesB = fixB ? es * (size_t)ldb * (size_t)g.rowsB : 0;
    char *ws = nullptr;
    ACC_CUDA(scratch_alloc(reinterpret_cast<void **>(&ws), bytesA + bytesB, g.stream));
    // one launch copies both operands (blockIdx.z): at the sizes where this path matters (250^3 Floats take
    // ~15 us in all) a second copy launch is a visible fraction of the call
    const size_t w = es < 8 ? es : 8;                                  // copy word: 4 bytes (f32) or 8
    RepitchArgs ra{}, rb{};
    if (fixA) {
        ra = RepitchArgs{static_cast<const char *>(g.A), ws, g.rowsA, (int64_t)(es * (size_t)g.colsA / w),
                         (int64_t)(es * (size_t)g.lda / w), (int64_t)(es * (size_t)lda / w)};
        h.A = ws;
    }
    if (fixB) {
        rb = RepitchArgs{static_cast<const char *>(g.B), ws + bytesA, g.rowsB, (int64_t)(es * (size_t)g.colsB / w),
                         (int64_t)(es * (size_t)g.ldb / w), (int64_t)(es * (size_t)ldb / w)};
        h.B = ws + bytesA;
    }
    const int64_t words = std::max(ra.words, rb.words), rows = std::max(ra.rows, rb.rows);
    dim3 grid((unsigned)((words + 255) / 256), (unsigned)std::min<int64_t>(rows, 32768), 2);
    if (w == 4) repitch_kernel<uint32_t><<<grid, 256, 0, g.stream>>>(ra, rb);
    else        repitch_kernel<uint64_t><<<grid, 256, 0, g.stream>>>(ra, rb);
    int st = check_launch("repitch_kernel");
    if (st == ACCBLAS_OK) st = run_family(pick_family(h), h);
    cudaError_t e2 = cudaFreeAsync(ws, g.stream);
    if (st == ACCBLAS_OK && e2 != cudaSuccess) return cuda_fail(e2, "cudaFreeAsync");
    return st;
}
}  // namespace

namespace {
// RAII for stream-ordered scratch used by the host-buffer entry points
struct Scratch {
    void *p = nullptr;
    cudaStream_t s;
    explicit Scratch(cudaStream_t st) : s(st) {}
    cudaError_t alloc(size_t bytes) { return scratch_alloc(&p, bytes, s); }
    ~Scratch() { if (p) cudaFreeAsync(p, s); }
};

}  // namespace
}  // namespace accblas

using namespace accblas;

extern "C" {

int accblas_version(void) { return ACCBLAS_B200_VERSION; }

const char *accblas_status_string(int s)
{
    switch (s) {
    case ACCBLAS_OK: return "ok";
    case ACCBLAS_ERR_DTYPE: return "unknown dtype";
    case ACCBLAS_ERR_OP: return "unknown transpose op";
    case ACCBLAS_ERR_SHAPE: return "shape mismatch";
    case ACCBLAS_ERR_NULL: return "null pointer with non-empty extent";
    case ACCBLAS_ERR_CUDA: return "CUDA error";
    case ACCBLAS_ERR_ARCH: return "device is not sm_100 (B200)";
    case ACCBLAS_ERR_WORKSPACE: return "workspace too small or misaligned";
    case ACCBLAS_ERR_ALIAS: return "output aliases an input";
    default: return "unknown status";
    }
}

const char *accblas_last_cuda_error(void) { return g_cuda_err; }
size_t accblas_last_cuda_error_r(char *buf, size_t cap)
{
    std::lock_guard<std::mutex> lock(g_err_mu);
    const size_t n = strlen(g_cuda_err_any);
    if (buf && cap) {
        const size_t c = n < cap - 1 ? n : cap - 1;
        memcpy(buf, g_cuda_err_any, c);
        buf[c] = 0;
    }
    return n;
}
uint64_t accblas_launch_count(void) { return g_launches; }
uint64_t accblas_launch_count_total(void) { return g_launches_total.load(std::memory_order_relaxed); }

int accblas_trim_workspace(void)
{
    int dev = 0;
    ACC_CUDA(cudaGetDevice(&dev));
    if (dev < 0 || dev >= 64) return ACCBLAS_OK;
    cudaMemPool_t pool = nullptr;
    {
        std::lock_guard<std::mutex> lock(g_pool_mu);
        pool = g_pools[dev];
    }
    if (pool) ACC_CUDA(cudaMemPoolTrimTo(pool, 0));
    return ACCBLAS_OK;
}

size_t accblas_workspace_cached_bytes(void)
{
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return 0;
    cudaMemPool_t pool = nullptr;
    {
        std::lock_guard<std::mutex> lock(g_pool_mu);
        pool = g_pools[dev];
    }
    uint64_t v = 0;
    if (pool && cudaMemPoolGetAttribute(pool, cudaMemPoolAttrReservedMemCurrent, &v) != cudaSuccess) v = 0;
    return (size_t)v;
}
void accblas_set_gemm_path(int path) { g_gemm_path = (path >= 0 && path <= 2) ? path : 0; }
void accblas_set_tuning(int knob, int value)
{
    if (knob >= 0 && knob < ACCBLAS_TUNE_COUNT) g_tuning[knob] = value;
}

int accblas_gemm(accblas_stream_t stream, int dtype, int opA, int opB, int64_t rowsA, int64_t colsA,
                 int64_t rowsB, int64_t colsB, const void *alpha_host, const void *A, const void *B, void *C)
{
    GemmArgs g{};
    g.dtype = dtype; g.opA = opA; g.opB = opB;
    g.rowsA = rowsA; g.colsA = colsA; g.rowsB = rowsB; g.colsB = colsB;
    g.alpha = alpha_host; g.A = A; g.B = B; g.C = C;
    g.stream = static_cast<cudaStream_t>(stream);
    return gemm_device(g);
}

int accblas_gemm_ex(accblas_stream_t stream, int dtype, int opA, int opB, int64_t rowsA, int64_t colsA,
                    int64_t rowsB, int64_t colsB, const void *alpha_host, const void *A, int64_t lda,
                    const void *B, int64_t ldb, const void *beta_host, const void *Cin, int64_t ldcin,
                    void *C, int64_t ldc)
{
    if (lda < 0 || ldb < 0 || ldc < 0 || ldcin < 0) return ACCBLAS_ERR_SHAPE;
    GemmArgs g{};
    g.dtype = dtype; g.opA = opA; g.opB = opB;
    g.rowsA = rowsA; g.colsA = colsA; g.rowsB = rowsB; g.colsB = colsB;
    g.lda = lda; g.ldb = ldb; g.ldc = ldc;
    g.alpha = alpha_host; g.A = A; g.B = B; g.C = C;
    g.beta = beta_host; g.Cin = Cin; g.ldcin = ldcin;
    g.stream = static_cast<cudaStream_t>(stream);
    if (bad_dtype(dtype)) return ACCBLAS_ERR_DTYPE;
    if (beta_host && !Cin && has_beta_value(dtype, beta_host)) return ACCBLAS_ERR_NULL;
    return gemm_device(g);
}

int accblas_gemm_ld(accblas_stream_t stream, int dtype, int opA, int opB, int64_t rowsA, int64_t colsA,
                    int64_t rowsB, int64_t colsB, const void *alpha_host, const void *A, int64_t lda,
                    const void *B, int64_t ldb, void *C, int64_t ldc)
{
    if (lda < 0 || ldb < 0 || ldc < 0) return ACCBLAS_ERR_SHAPE;
    GemmArgs g{};
    g.dtype = dtype; g.opA = opA; g.opB = opB;
    g.rowsA = rowsA; g.colsA = colsA; g.rowsB = rowsB; g.colsB = colsB;
    g.lda = lda; g.ldb = ldb; g.ldc = ldc;
    g.alpha = alpha_host; g.A = A; g.B = B; g.C = C;
    g.stream = static_cast<cudaStream_t>(stream);
    return gemm_device(g);
}

// Strided batch of independent products with identical shapes and flags.  Small products (the ones the SIMT
// kernel serves) run as ONE launch with the batch on gridDim.z; products a tensor family takes are enqueued one
// after the other on the stream (each already fills the GPU).
int accblas_gemm_strided_batched(accblas_stream_t stream, int dtype, int opA, int opB, int64_t rowsA, int64_t colsA,
                                 int64_t rowsB, int64_t colsB, const void *alpha_host, const void *A, int64_t strideA,
                                 const void *B, int64_t strideB, void *C, int64_t strideC, int64_t batch)
{
    if (batch < 0 || strideA < 0 || strideB < 0 || strideC < 0) return ACCBLAS_ERR_SHAPE;
    GemmArgs g{};
    g.dtype = dtype; g.opA = opA; g.opB = opB;
    g.rowsA = rowsA; g.colsA = colsA; g.rowsB = rowsB; g.colsB = colsB;
    g.alpha = alpha_host; g.A = A; g.B = B; g.C = C;
    g.stream = static_cast<cudaStream_t>(stream);
    ACC_TRY(derive_gemm(g));
    if (batch == 0 || g.m == 0 || g.n == 0) return ACCBLAS_OK;
    if (batch > 1 && strideC < g.m * g.n) return ACCBLAS_ERR_ALIAS;           // outputs of the batch must not overlap
    if (g.k > 0 && ((strideA != 0 && strideA < g.rowsA * g.colsA) || (strideB != 0 && strideB < g.rowsB * g.colsB)))
        return ACCBLAS_ERR_SHAPE;
    const size_t es = elem_size(dtype);
    auto at = [&](const void *p, int64_t stride, int64_t b) {
        return p ? static_cast<const void *>(static_cast<const char *>(p) + es * (size_t)stride * (size_t)b) : p;
    };
    if (batch > 1 && g.k > 0 && g.A && g.B && g.C && g.alpha && g_gemm_path != 2 && pick_family(g) == FAM_GENERIC &&
        (double)g.m * (double)g.n * (double)g.k < 48.0 * 48.0 * 48.0) {
        // every product of the batch is tiny: one z-batched launch per 65535 products
        ACC_TRY(device_check());
        const size_t spanC = es * ((size_t)strideC * (size_t)(batch - 1) + (size_t)g.m * (size_t)g.n);
        if (overlaps(g.C, spanC, g.A, es * ((size_t)strideA * (size_t)(batch - 1) + span(g.rowsA, g.colsA, g.lda))) ||
            overlaps(g.C, spanC, g.B, es * ((size_t)strideB * (size_t)(batch - 1) + span(g.rowsB, g.colsB, g.ldb))))
            return ACCBLAS_ERR_ALIAS;
        for (int64_t b0 = 0; b0 < batch; b0 += 65535) {
            GemmArgs h = g;
            h.batch = std::min<int64_t>(65535, batch - b0);
            h.strideA = strideA; h.strideB = strideB; h.strideC = strideC;
            h.A = at(A, strideA, b0); h.B = at(B, strideB, b0); h.C = const_cast<void *>(at(C, strideC, b0));
            ACC_TRY(gemm_generic(h));
        }
        return ACCBLAS_OK;
    }
    // products of the warp-level tensor families (DMMA: f64 / c64, mma.sync: f32 / c32 -- everything below the sizes that
    // fill the GPU on their own): ONE launch, the batch folded into the grid (blockIdx -> (product, tile)) and into a third
    // dimension of the TMA tensor maps, so the tiles of all products share the waves.  Needs 16-byte multiples for the
    // strides (TMA plane stride, vector stores); anything else falls through to the loop of launches below.
    if (batch > 1 && g.k > 0 && g.A && g.B && g.C && g.alpha && g_gemm_path == 0) {
        const Family f = pick_family(g);
        const auto ok16 = [&](int64_t stride) { return (es * (size_t)stride) % 16 == 0; };
        if ((f == FAM_DMMA || f == FAM_MMA) && ok16(strideA) && ok16(strideB) && ok16(strideC) && batch <= 0x7fffffff) {
            ACC_TRY(device_check());
            const size_t spanC = es * ((size_t)strideC * (size_t)(batch - 1) + (size_t)g.m * (size_t)g.n);
            if (overlaps(g.C, spanC, g.A, es * ((size_t)strideA * (size_t)(batch - 1) + span(g.rowsA, g.colsA, g.lda))) ||
                overlaps(g.C, spanC, g.B, es * ((size_t)strideB * (size_t)(batch - 1) + span(g.rowsB, g.colsB, g.ldb))))
                return ACCBLAS_ERR_ALIAS;
            GemmArgs h = g;
            h.batch = batch; h.strideA = strideA; h.strideB = strideB; h.strideC = strideC;
            h.no_persistent = true;
            return run_family(f, h);
        }
    }
    for (int64_t b = 0; b < batch; ++b) {
        GemmArgs h{};
        h.dtype = dtype; h.opA = opA; h.opB = opB;
        h.rowsA = rowsA; h.colsA = colsA; h.rowsB = rowsB; h.colsB = colsB;
        h.alpha = alpha_host; h.stream = g.stream;
        h.A = at(A, strideA, b); h.B = at(B, strideB, b); h.C = const_cast<void *>(at(C, strideC, b));
        ACC_TRY(gemm_device(h));
    }
    return ACCBLAS_OK;
}

const char *accblas_gemm_kernel_name(int dtype, int opA, int opB, int64_t rowsA, int64_t colsA,
                                     int64_t rowsB, int64_t colsB)
{
    GemmArgs g{};
    g.dtype = dtype; g.opA = opA; g.opB = opB;
    g.rowsA = rowsA; g.colsA = colsA; g.rowsB = rowsB; g.colsB = colsB;
    if (derive_gemm(g) != ACCBLAS_OK) return "invalid";
    Family f = pick_family(g);
    if (f == FAM_GENERIC && g_gemm_path == 0 && (double)g.m * (double)g.n * (double)g.k >= 48.0 * 48.0 * 48.0) {
        if (g_tuning[ACCBLAS_TUNE_SGEMM_DIRECT] != 0 && sgemm_mma_direct_eligible(g)) return "sgemm_tf32x3_mma_direct";
        if (g.dtype == ACCBLAS_F64 && dgemm_direct_eligible(g)) return "dgemm_dmma_direct";
        const int64_t q = (int64_t)(16 / elem_size(g.dtype));         // what gemm_repitched would run
        GemmArgs h = g;
        h.lda = (g.colsA + q - 1) / q * q;
        h.ldb = (g.colsB + q - 1) / q * q;
        f = pick_family(h);
    }
    switch (f) {
    case FAM_NONE: return "none";
    case FAM_DMMA: return g.dtype == ACCBLAS_F64 ? "dgemm_dmma_tma" : "zgemm_dmma_tma";
    case FAM_TC:   return g.dtype == ACCBLAS_F32 ? "sgemm_tf32x3_tcgen05" : "cgemm_tf32x3_tcgen05";
    case FAM_RANK1: return "gemm_rank1_outer";
    case FAM_MMA:  return g.dtype == ACCBLAS_F32 ? "sgemm_tf32x3_mma" : "cgemm_tf32x3_mma";
    case FAM_Z3M:  return "zgemm_3m_dmma";
    case FAM_C3M:  return "cgemm_3m_tf32x3_tcgen05";
    default:       return "gemm_simt_generic";
    }
}

size_t accblas_gemm_workspace_bytes(int dtype, int opA, int opB, int64_t rowsA, int64_t colsA,
                                    int64_t rowsB, int64_t colsB)
{
    GemmArgs g{};
    g.dtype = dtype; g.opA = opA; g.opB = opB;
    g.rowsA = rowsA; g.colsA = colsA; g.rowsB = rowsB; g.colsB = colsB;
    if (derive_gemm(g) != ACCBLAS_OK) return 0;
    const Family f = pick_family(g);
    return f == FAM_TC ? sgemm_tc_workspace(g) : f == FAM_Z3M ? zgemm_3m_workspace(g) : f == FAM_C3M ? cgemm_3m_workspace(g)
         : f == FAM_DMMA ? dgemm_dmma_workspace(g) : 0;
}

static int gemv_entry(accblas_stream_t stream, int dtype, int opA, int64_t rowsA, int64_t colsA,
                      const void *alpha_host, bool alpha_dev, const void *A, const void *x, int64_t lenx, void *y)
{
    if (bad_dtype(dtype)) return ACCBLAS_ERR_DTYPE;
    if (bad_op(opA)) return ACCBLAS_ERR_OP;
    if (rowsA < 0 || colsA < 0 || lenx < 0) return ACCBLAS_ERR_SHAPE;
    const int64_t need = opA == ACCBLAS_OP_N ? colsA : rowsA;
    const int64_t leny = opA == ACCBLAS_OP_N ? rowsA : colsA;   // LLVM/PTX/Level2.hs:80-82
    if (lenx != need) return ACCBLAS_ERR_SHAPE;
    if (leny == 0) return ACCBLAS_OK;
    if (!y || !alpha_host) return ACCBLAS_ERR_NULL;
    if (need > 0 && (!A || !x)) return ACCBLAS_ERR_NULL;
    const size_t es = elem_size(dtype);
    if (overlaps(y, es * (size_t)leny, A, es * (size_t)rowsA * (size_t)colsA) ||
        overlaps(y, es * (size_t)leny, x, es * (size_t)lenx))
        return ACCBLAS_ERR_ALIAS;
    ACC_TRY(device_check());
    if (!is_complex(dtype) && opA == ACCBLAS_OP_H) opA = ACCBLAS_OP_T;
    return gemv_dispatch(static_cast<cudaStream_t>(stream), dtype, opA, rowsA, colsA, alpha_host, alpha_dev, A, x, y);
}

int accblas_gemv(accblas_stream_t stream, int dtype, int opA, int64_t rowsA, int64_t colsA,
                 const void *alpha_host, const void *A, const void *x, int64_t lenx, void *y)
{
    return gemv_entry(stream, dtype, opA, rowsA, colsA, alpha_host, false, A, x, lenx, y);
}

int accblas_gemv_alpha_dev(accblas_stream_t stream, int dtype, int opA, int64_t rowsA, int64_t colsA,
                           const void *alpha_dev, const void *A, const void *x, int64_t lenx, void *y)
{
    return gemv_entry(stream, dtype, opA, rowsA, colsA, alpha_dev, true, A, x, lenx, y);
}

int accblas_gemm_alpha_dev(accblas_stream_t stream, int dtype, int opA, int opB, int64_t rowsA, int64_t colsA,
                           int64_t rowsB, int64_t colsB, const void *alpha_dev, const void *A, const void *B, void *C)
{
    GemmArgs g{};
    g.dtype = dtype; g.opA = opA; g.opB = opB;
    g.rowsA = rowsA; g.colsA = colsA; g.rowsB = rowsB; g.colsB = colsB;
    g.alpha = alpha_dev; g.alpha_dev = true; g.A = A; g.B = B; g.C = C;
    g.stream = static_cast<cudaStream_t>(stream);
    return gemm_device(g);
}

int accblas_dot(accblas_stream_t stream, int dtype, int conj_x, int64_t n, const void *x, const void *y,
                void *out_dev)
{
    if (bad_dtype(dtype)) return ACCBLAS_ERR_DTYPE;
    if (n < 0) return ACCBLAS_ERR_SHAPE;
    if (!out_dev) return ACCBLAS_ERR_NULL;
    if (n > 0 && (!x || !y)) return ACCBLAS_ERR_NULL;
    ACC_TRY(device_check());
    return dot_dispatch(static_cast<cudaStream_t>(stream), dtype, conj_x, n, x, y, out_dev);
}

int accblas_dot_compensated(accblas_stream_t stream, int dtype, int conj_x, int64_t n, const void *x, const void *y,
                            void *out_dev)
{
    if (bad_dtype(dtype)) return ACCBLAS_ERR_DTYPE;
    if (n < 0) return ACCBLAS_ERR_SHAPE;
    if (!out_dev) return ACCBLAS_ERR_NULL;
    if (n > 0 && (!x || !y)) return ACCBLAS_ERR_NULL;
    ACC_TRY(device_check());
    return dot_dispatch(static_cast<cudaStream_t>(stream), dtype, conj_x, n, x, y, out_dev, true);
}

int accblas_sdot(accblas_stream_t stream, int dtype, int64_t n, const void *z_host, const void *x,
                 const void *y, void *out_dev)
{
    if (bad_dtype(dtype)) return ACCBLAS_ERR_DTYPE;
    if (n < 0) return ACCBLAS_ERR_SHAPE;
    if (!out_dev || !z_host) return ACCBLAS_ERR_NULL;
    if (n > 0 && (!x || !y)) return ACCBLAS_ERR_NULL;
    ACC_TRY(device_check());
    return sdot_dispatch(static_cast<cudaStream_t>(stream), dtype, n, z_host, x, y, out_dev);
}

int accblas_asum(accblas_stream_t stream, int dtype, int64_t n, const void *x, void *out_dev)
{
    if (bad_dtype(dtype)) return ACCBLAS_ERR_DTYPE;
    if (n < 0) return ACCBLAS_ERR_SHAPE;
    if (!out_dev || (n > 0 && !x)) return ACCBLAS_ERR_NULL;
    ACC_TRY(device_check());
    return asum_dispatch(static_cast<cudaStream_t>(stream), dtype, n, x, out_dev);
}

static int iamax_entry(accblas_stream_t stream, int dtype, int want_max, int64_t n, const void *x, int64_t *out_dev)
{
    if (bad_dtype(dtype)) return ACCBLAS_ERR_DTYPE;
    if (n <= 0) return ACCBLAS_ERR_SHAPE;           // fold1 over an empty vector is an error in the reference too
    if (!out_dev || !x) return ACCBLAS_ERR_NULL;
    ACC_TRY(device_check());
    return iamax_dispatch(static_cast<cudaStream_t>(stream), dtype, want_max, n, x, out_dev);
}
int accblas_amax(accblas_stream_t stream, int dtype, int64_t n, const void *x, int64_t *out_dev)
{
    return iamax_entry(stream, dtype, 1, n, x, out_dev);
}
int accblas_amin(accblas_stream_t stream, int dtype, int64_t n, const void *x, int64_t *out_dev)
{
    return iamax_entry(stream, dtype, 0, n, x, out_dev);
}

// -------------------------------------------------------------------------------------
// host-buffer entry points (upload -> kernels -> download, all on `stream`)
// -------------------------------------------------------------------------------------
namespace {
// the copy stream and events of the chunked gemv upload: per thread and device, created on first use and kept (like the side
// streams of host_pipeline.cu); they carry no state between calls -- the entry points are synchronous
struct HostPipe {
    cudaStream_t copy = nullptr;
    std::vector<cudaEvent_t> events;
    size_t used = 0;
    cudaError_t create() { return copy ? cudaSuccess : cudaStreamCreateWithFlags(&copy, cudaStreamNonBlocking); }
    cudaEvent_t event() {
        if (used == events.size()) {
            cudaEvent_t e = nullptr;
            cudaEventCreateWithFlags(&e, cudaEventDisableTiming);
            events.push_back(e);
        }
        return events[used++];
    }
};
HostPipe &host_pipe()
{
    static thread_local HostPipe per_device[64];
    int dev = 0;
    cudaGetDevice(&dev);
    HostPipe &h = per_device[(dev >= 0 && dev < 64) ? dev : 0];
    h.used = 0;
    return h;
}
}  // namespace
int accblas_gemm_host(accblas_stream_t stream, int dtype, int opA, int opB, int64_t rowsA, int64_t colsA,
                      int64_t rowsB, int64_t colsB, const void *alpha_host, const void *A_host,
                      const void *B_host, void *C_host)
{
    GemmArgs g{};
    g.dtype = dtype; g.opA = opA; g.opB = opB;
    g.rowsA = rowsA; g.colsA = colsA; g.rowsB = rowsB; g.colsB = colsB;
    g.alpha = alpha_host;
    g.stream = static_cast<cudaStream_t>(stream);
    ACC_TRY(derive_gemm(g));
    if (g.m == 0 || g.n == 0) return ACCBLAS_OK;
    if (!C_host || !alpha_host) return ACCBLAS_ERR_NULL;
    if (g.k > 0 && (!A_host || !B_host)) return ACCBLAS_ERR_NULL;
    ACC_TRY(device_check());
    return gemm_host_pipeline(g, A_host, B_host, C_host);
}

int accblas_gemv_host(accblas_stream_t stream, int dtype, int opA, int64_t rowsA, int64_t colsA,
                      const void *alpha_host, const void *A_host, const void *x_host, int64_t lenx,
                      void *y_host)
{
    if (bad_dtype(dtype)) return ACCBLAS_ERR_DTYPE;
    if (bad_op(opA)) return ACCBLAS_ERR_OP;
    if (rowsA < 0 || colsA < 0 || lenx < 0) return ACCBLAS_ERR_SHAPE;
    const int64_t need = opA == ACCBLAS_OP_N ? colsA : rowsA;
    const int64_t leny = opA == ACCBLAS_OP_N ? rowsA : colsA;
    if (lenx != need) return ACCBLAS_ERR_SHAPE;
    if (leny == 0) return ACCBLAS_OK;
    if (!y_host || !alpha_host || (need > 0 && (!A_host || !x_host))) return ACCBLAS_ERR_NULL;
    const size_t es = elem_size(dtype);
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    Scratch dA(s), dx(s), dy(s);
    ACC_CUDA(dA.alloc(es * (size_t)rowsA * (size_t)colsA));
    ACC_CUDA(dx.alloc(es * (size_t)lenx));
    ACC_CUDA(dy.alloc(es * (size_t)leny));
    // op N on a large matrix: row chunks of A go up on a second stream and chunk i multiplies (y[chunk] = alpha * A[chunk, :] x)
    // while chunk i + 1 is on the link -- the product is < 1 % of the transfer (1.2 ms against 155 ms for 32768^2 doubles),
    // so all this hides is the product.  op T / H needs every row of A for every output and stays upload -> product -> download.
    const size_t row_bytes = es * (size_t)colsA;
    const int64_t chunk_rows = row_bytes ? std::max<int64_t>(1, (int64_t)(((size_t)16 << 20) / row_bytes)) : rowsA;
    if (need > 0 && opA == ACCBLAS_OP_N && rowsA >= 4 * chunk_rows) {
        HostPipe &hp = host_pipe();
        ACC_CUDA(hp.create());
        ACC_CUDA(cudaMemcpyAsync(dx.p, x_host, es * (size_t)lenx, cudaMemcpyHostToDevice, s));
        cudaEvent_t start = hp.event();
        ACC_CUDA(cudaEventRecord(start, s));
        ACC_CUDA(cudaStreamWaitEvent(hp.copy, start, 0));
        int st = ACCBLAS_OK;
        cudaError_t ce = cudaSuccess;                 // no early return inside the loop: both streams are joined before the scratch goes
        for (int64_t r0 = 0; r0 < rowsA && st == ACCBLAS_OK && ce == cudaSuccess; r0 += chunk_rows) {
            const int64_t h = std::min(chunk_rows, rowsA - r0);
            char *dchunk = static_cast<char *>(dA.p) + row_bytes * (size_t)r0;
            ce = cudaMemcpyAsync(dchunk, static_cast<const char *>(A_host) + row_bytes * (size_t)r0, row_bytes * (size_t)h,
                                 cudaMemcpyHostToDevice, hp.copy);
            cudaEvent_t landed = hp.event();
            if (ce == cudaSuccess) ce = cudaEventRecord(landed, hp.copy);
            if (ce == cudaSuccess) ce = cudaStreamWaitEvent(s, landed, 0);
            if (ce == cudaSuccess)
                st = accblas_gemv(stream, dtype, opA, h, colsA, alpha_host, dchunk, dx.p, lenx, static_cast<char *>(dy.p) + es * (size_t)r0);
        }
        if (st == ACCBLAS_OK && ce == cudaSuccess) ce = cudaMemcpyAsync(y_host, dy.p, es * (size_t)leny, cudaMemcpyDeviceToHost, s);
        cudaStreamSynchronize(hp.copy);
        const cudaError_t se = cudaStreamSynchronize(s);
        if (st != ACCBLAS_OK) return st;
        if (ce != cudaSuccess) return cuda_fail(ce, "accblas_gemv_host (chunked upload)");
        if (se != cudaSuccess) return cuda_fail(se, "cudaStreamSynchronize");
        return ACCBLAS_OK;
    }
    if (need > 0) {
        ACC_CUDA(cudaMemcpyAsync(dA.p, A_host, es * (size_t)rowsA * (size_t)colsA, cudaMemcpyHostToDevice, s));
        ACC_CUDA(cudaMemcpyAsync(dx.p, x_host, es * (size_t)lenx, cudaMemcpyHostToDevice, s));
    }
    ACC_TRY(accblas_gemv(stream, dtype, opA, rowsA, colsA, alpha_host, dA.p, dx.p, lenx, dy.p));
    ACC_CUDA(cudaMemcpyAsync(y_host, dy.p, es * (size_t)leny, cudaMemcpyDeviceToHost, s));
    ACC_CUDA(cudaStreamSynchronize(s));
    return ACCBLAS_OK;
}

int accblas_dot_host(accblas_stream_t stream, int dtype, int conj_x, int64_t n, const void *x_host,
                     const void *y_host, void *out_host)
{
    if (bad_dtype(dtype)) return ACCBLAS_ERR_DTYPE;
    if (n < 0) return ACCBLAS_ERR_SHAPE;
    if (!out_host || (n > 0 && (!x_host || !y_host))) return ACCBLAS_ERR_NULL;
    const size_t es = elem_size(dtype);
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    Scratch dx(s), dy(s), dout(s);
    ACC_CUDA(dx.alloc(es * (size_t)n));
    ACC_CUDA(dy.alloc(es * (size_t)n));
    ACC_CUDA(dout.alloc(es));
    if (n > 0) {
        ACC_CUDA(cudaMemcpyAsync(dx.p, x_host, es * (size_t)n, cudaMemcpyHostToDevice, s));
        ACC_CUDA(cudaMemcpyAsync(dy.p, y_host, es * (size_t)n, cudaMemcpyHostToDevice, s));
    }
    ACC_TRY(accblas_dot(stream, dtype, conj_x, n, dx.p, dy.p, dout.p));
    ACC_CUDA(cudaMemcpyAsync(out_host, dout.p, es, cudaMemcpyDeviceToHost, s));
    ACC_CUDA(cudaStreamSynchronize(s));
    return ACCBLAS_OK;
}

// One-vector / two-vector Level-1 host wrappers: upload, reduce, download the scalar.
static int level1_host(cudaStream_t s, size_t es, int64_t n, const void *x_host, const void *y_host,
                       size_t out_bytes, void *out_host, int (*run)(cudaStream_t, const void *, const void *, void *, void *),
                       void *ctx)
{
    Scratch dx(s), dy(s), dout(s);
    ACC_CUDA(dx.alloc(es * (size_t)n));
    if (y_host) ACC_CUDA(dy.alloc(es * (size_t)n));
    ACC_CUDA(dout.alloc(out_bytes));
    if (n > 0) {
        ACC_CUDA(cudaMemcpyAsync(dx.p, x_host, es * (size_t)n, cudaMemcpyHostToDevice, s));
        if (y_host) ACC_CUDA(cudaMemcpyAsync(dy.p, y_host, es * (size_t)n, cudaMemcpyHostToDevice, s));
    }
    ACC_TRY(run(s, dx.p, y_host ? dy.p : nullptr, dout.p, ctx));
    ACC_CUDA(cudaMemcpyAsync(out_host, dout.p, out_bytes, cudaMemcpyDeviceToHost, s));
    ACC_CUDA(cudaStreamSynchronize(s));
    return ACCBLAS_OK;
}

struct L1Ctx { int dtype; int64_t n; const void *z; int want_max; };

int accblas_sdot_host(accblas_stream_t stream, int dtype, int64_t n, const void *z_host, const void *x_host,
                      const void *y_host, void *out_host)
{
    if (bad_dtype(dtype)) return ACCBLAS_ERR_DTYPE;
    if (n < 0) return ACCBLAS_ERR_SHAPE;
    if (!out_host || !z_host || (n > 0 && (!x_host || !y_host))) return ACCBLAS_ERR_NULL;
    L1Ctx c{dtype, n, z_host, 0};
    static const char dummy = 0;
    return level1_host(static_cast<cudaStream_t>(stream), elem_size(dtype), n, x_host, y_host ? y_host : &dummy,
                       elem_size(dtype), out_host,
                       [](cudaStream_t s, const void *x, const void *y, void *o, void *cv) {
                           auto *c = static_cast<L1Ctx *>(cv);
                           return accblas_sdot(s, c->dtype, c->n, c->z, x, y, o);
                       }, &c);
}

int accblas_asum_host(accblas_stream_t stream, int dtype, int64_t n, const void *x_host, void *out_host)
{
    if (bad_dtype(dtype)) return ACCBLAS_ERR_DTYPE;
    if (n < 0) return ACCBLAS_ERR_SHAPE;
    if (!out_host || (n > 0 && !x_host)) return ACCBLAS_ERR_NULL;
    L1Ctx c{dtype, n, nullptr, 0};
    const size_t ob = (dtype == ACCBLAS_F32 || dtype == ACCBLAS_C32) ? 4 : 8;
    return level1_host(static_cast<cudaStream_t>(stream), elem_size(dtype), n, x_host, nullptr, ob, out_host,
                       [](cudaStream_t s, const void *x, const void *, void *o, void *cv) {
                           auto *c = static_cast<L1Ctx *>(cv);
                           return accblas_asum(s, c->dtype, c->n, x, o);
                       }, &c);
}

static int iamax_host(accblas_stream_t stream, int dtype, int want_max, int64_t n, const void *x_host, int64_t *out_host)
{
    if (bad_dtype(dtype)) return ACCBLAS_ERR_DTYPE;
    if (n <= 0) return ACCBLAS_ERR_SHAPE;
    if (!out_host || !x_host) return ACCBLAS_ERR_NULL;
    L1Ctx c{dtype, n, nullptr, want_max};
    return level1_host(static_cast<cudaStream_t>(stream), elem_size(dtype), n, x_host, nullptr, 8, out_host,
                       [](cudaStream_t s, const void *x, const void *, void *o, void *cv) {
                           auto *c = static_cast<L1Ctx *>(cv);
                           return iamax_entry(s, c->dtype, c->want_max, c->n, x, static_cast<int64_t *>(o));
                       }, &c);
}
int accblas_amax_host(accblas_stream_t stream, int dtype, int64_t n, const void *x_host, int64_t *out_host)
{
    return iamax_host(stream, dtype, 1, n, x_host, out_host);
}
int accblas_amin_host(accblas_stream_t stream, int dtype, int64_t n, const void *x_host, int64_t *out_host)
{
    return iamax_host(stream, dtype, 0, n, x_host, out_host);
}

}  // extern "C"
