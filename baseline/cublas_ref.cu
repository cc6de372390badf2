// baseline/cublas_ref.cu -- REPORTED BASELINE ONLY (never linked into libaccblas_b200.so).
//
// The reference's llvm-ptx path restated call for call: one cublas?gemm / cublas?gemv per product
// with the row-major operand swap, host pointer mode, beta = 0, default math mode, atomics allowed
//   BLAS.dgemm hdl opB' opA' n m k alpha B ldb A lda beta C n      LLVM/PTX/Level3.hs:96
//   BLAS.dgemv hdl op'  colsA rowsA alpha A colsA x 1 beta y 1     LLVM/PTX/Level2.hs:114
//   setAtomicsMode hdl AtomicsAllowed                              LLVM/PTX/Context.hs:59
// cublasDdot is a labelled stand-in: the reference's dot is Accelerate's own fold (Level1.hs:74).
// Prints one JSON line per case.  usage: cublas_ref [case-substring] [iters]
#include <cublas_v2.h>
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>
#include <algorithm>

#define CK(x) do { auto e_ = (x); if (e_ != 0) { fprintf(stderr, "fail %s -> %d line %d\n", #x, (int)e_, __LINE__); exit(1); } } while (0)

__global__ void fill(double *p, size_t n, unsigned seed) {   // uniform(-1,1), cheap LCG hash
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
        unsigned long long z = (i + 1) * 0x9E3779B97F4A7C15ull + seed;
        z ^= z >> 30; z *= 0xBF58476D1CE4E5B9ull; z ^= z >> 27; z *= 0x94D049BB133111EBull; z ^= z >> 31;
        p[i] = (double)(z >> 11) * (2.0 / 9007199254740992.0) - 1.0;
    }
}
__global__ void fillf(float *p, size_t n, unsigned seed) {
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
        unsigned long long z = (i + 1) * 0x9E3779B97F4A7C15ull + seed;
        z ^= z >> 30; z *= 0xBF58476D1CE4E5B9ull; z ^= z >> 27; z *= 0x94D049BB133111EBull; z ^= z >> 31;
        p[i] = (float)((double)(z >> 11) * (2.0 / 9007199254740992.0) - 1.0);
    }
}

static cublasHandle_t H;
static const char *filter = "";
static int iters_override = 0;

// the filter is a comma-separated list of substrings: a case runs if any of them occurs in its name
static bool wanted(const char *name) {
    if (!*filter) return true;
    const char *f = filter;
    while (*f) {
        const char *e = strchr(f, ',');
        size_t len = e ? (size_t)(e - f) : strlen(f);
        if (len > 0 && len < 64) {
            char tok[64];
            memcpy(tok, f, len); tok[len] = 0;
            if (strstr(name, tok)) return true;
        }
        if (!e) break;
        f = e + 1;
    }
    return false;
}

// Calls are enqueued back to back with an event between them (the same method as tools/perf.py): the GPU
// never idles between launches, so no launch latency is billed to the kernel.
template <typename F> float time_ms(F f, int warm, int iters, float *best) {
    std::vector<cudaEvent_t> ev(iters + 1);
    for (auto &e : ev) cudaEventCreate(&e);
    for (int i = 0; i < warm; ++i) f();
    CK(cudaDeviceSynchronize());
    cudaEventRecord(ev[0]);
    for (int i = 0; i < iters; ++i) { f(); cudaEventRecord(ev[i + 1]); }
    CK(cudaDeviceSynchronize());
    std::vector<float> t(iters);
    for (int i = 0; i < iters; ++i) cudaEventElapsedTime(&t[i], ev[i], ev[i + 1]);
    for (auto &e : ev) cudaEventDestroy(e);
    std::sort(t.begin(), t.end());
    *best = t[0];
    return t[iters / 2];
}

static cublasOperation_t enc(char c) { return c == 'N' ? CUBLAS_OP_N : c == 'T' ? CUBLAS_OP_T : CUBLAS_OP_C; }

// dtype: s d c z ; reference argument order (PTX/Level3.hs:91-106)
static void gemm_case(char dt, char opA, char opB, long m, long n, long k, int iters) {
    char name[64]; snprintf(name, sizeof name, "%cgemm_%c%c_%ldx%ldx%ld", dt, opA, opB, m, n, k);
    if (!wanted(name)) return;
    if (iters_override) iters = iters_override;
    size_t es = dt == 's' ? 4 : dt == 'z' ? 16 : 8;
    void *A, *B, *C;
    CK(cudaMalloc(&A, es * m * k)); CK(cudaMalloc(&B, es * k * n)); CK(cudaMalloc(&C, es * m * n));
    if (dt == 's' || dt == 'c') { fillf<<<1184, 256>>>((float *)A, es / 4 * m * k, 1); fillf<<<1184, 256>>>((float *)B, es / 4 * k * n, 2); }
    else { fill<<<1184, 256>>>((double *)A, es / 8 * m * k, 1); fill<<<1184, 256>>>((double *)B, es / 8 * k * n, 2); }
    long rowsA = opA == 'N' ? m : k, colsA = opA == 'N' ? k : m, colsB = opB == 'N' ? n : k; (void)rowsA;
    int lda = (int)colsA, ldb = (int)colsB;
    auto call = [&]() {
        if (dt == 's') { float al = 1, be = 0; CK(cublasSgemm(H, enc(opB), enc(opA), n, m, k, &al, (float *)B, ldb, (float *)A, lda, &be, (float *)C, n)); }
        if (dt == 'd') { double al = 1, be = 0; CK(cublasDgemm(H, enc(opB), enc(opA), n, m, k, &al, (double *)B, ldb, (double *)A, lda, &be, (double *)C, n)); }
        if (dt == 'c') { cuComplex al = {1, 0}, be = {0, 0}; CK(cublasCgemm(H, enc(opB), enc(opA), n, m, k, &al, (cuComplex *)B, ldb, (cuComplex *)A, lda, &be, (cuComplex *)C, n)); }
        if (dt == 'z') { cuDoubleComplex al = {1, 0}, be = {0, 0}; CK(cublasZgemm(H, enc(opB), enc(opA), n, m, k, &al, (cuDoubleComplex *)B, ldb, (cuDoubleComplex *)A, lda, &be, (cuDoubleComplex *)C, n)); }
    };
    float best, med = time_ms(call, m >= 16384 ? 1 : 3, iters, &best);
    double flop = ((dt == 'c' || dt == 'z') ? 8.0 : 2.0) * m * n * k;
    printf("{\"impl\":\"cublas\",\"case\":\"%s\",\"ms_median\":%.4f,\"ms_best\":%.4f,\"tflops_median\":%.3f,\"tflops_best\":%.3f}\n",
           name, med, best, flop / med * 1e-9, flop / best * 1e-9);
    fflush(stdout);
    cudaFree(A); cudaFree(B); cudaFree(C);
}

static void gemv_case(char op, long rows, long cols, int iters) {
    char name[64]; snprintf(name, sizeof name, "dgemv_%c_%ldx%ld", op, rows, cols);
    if (!wanted(name)) return;
    if (iters_override) iters = iters_override;
    double *A, *x, *y;
    long lenx = op == 'N' ? cols : rows, leny = op == 'N' ? rows : cols;
    CK(cudaMalloc(&A, 8 * rows * cols)); CK(cudaMalloc(&x, 8 * lenx)); CK(cudaMalloc(&y, 8 * leny));
    fill<<<1184, 256>>>(A, rows * cols, 3); fill<<<64, 256>>>(x, lenx, 4);
    cublasOperation_t o = op == 'N' ? CUBLAS_OP_T : CUBLAS_OP_N;      // PTX/Level2.hs:84-87
    auto call = [&]() { double al = 1, be = 0; CK(cublasDgemv(H, o, (int)cols, (int)rows, &al, A, (int)cols, x, 1, &be, y, 1)); };
    float best, med = time_ms(call, 3, iters, &best);
    double bytes = 8.0 * (rows * cols + lenx + leny);
    printf("{\"impl\":\"cublas\",\"case\":\"%s\",\"ms_median\":%.4f,\"ms_best\":%.4f,\"gbs_median\":%.1f,\"gbs_best\":%.1f}\n",
           name, med, best, bytes / med * 1e-6, bytes / best * 1e-6);
    fflush(stdout);
    cudaFree(A); cudaFree(x); cudaFree(y);
}

static void dot_case(long n, int iters) {
    char name[64]; snprintf(name, sizeof name, "ddot_%ld", n);
    if (!wanted(name)) return;
    if (iters_override) iters = iters_override;
    double *x, *y, *out;
    CK(cudaMalloc(&x, 8 * n)); CK(cudaMalloc(&y, 8 * n)); CK(cudaMalloc(&out, 8));
    fill<<<1184, 256>>>(x, n, 5); fill<<<1184, 256>>>(y, n, 6);
    CK(cublasSetPointerMode(H, CUBLAS_POINTER_MODE_DEVICE));          // result stays on the device
    auto call = [&]() { CK(cublasDdot(H, (int)n, x, 1, y, 1, out)); };
    float best, med = time_ms(call, 3, iters, &best);
    CK(cublasSetPointerMode(H, CUBLAS_POINTER_MODE_HOST));
    double bytes = 16.0 * n + 8;
    printf("{\"impl\":\"cublas\",\"case\":\"%s\",\"note\":\"stand-in: reference dot is Accelerate fold\",\"ms_median\":%.4f,\"ms_best\":%.4f,\"gbs_median\":%.1f,\"gbs_best\":%.1f}\n",
           name, med, best, bytes / med * 1e-6, bytes / best * 1e-6);
    fflush(stdout);
    cudaFree(x); cudaFree(y); cudaFree(out);
}

int main(int argc, char **argv) {
    if (argc > 1) filter = argv[1];
    if (argc > 2) iters_override = atoi(argv[2]);
    CK(cublasCreate(&H));
    CK(cublasSetAtomicsMode(H, CUBLAS_ATOMICS_ALLOWED));              // PTX/Context.hs:59
    gemm_case('d', 'N', 'N', 1024, 1024, 1024, 50);
    gemm_case('d', 'N', 'N', 1000, 1000, 1000, 50);
    for (long n : {100L, 250L, 500L, 512L, 2048L, 4096L}) gemm_case('d', 'N', 'N', n, n, n, 50);
    for (long n : {100L, 250L, 500L, 1000L}) gemm_case('s', 'N', 'N', n, n, n, 50);
    for (const char *ops : {"NN", "NT", "TN"}) {
        gemm_case('d', ops[0], ops[1], 8192, 8192, 8192, 10);
        gemm_case('s', ops[0], ops[1], 8192, 8192, 8192, 10);
    }
    gemm_case('z', 'N', 'N', 4096, 4096, 4096, 10);
    gemm_case('z', 'H', 'N', 4096, 4096, 4096, 10);
    gemm_case('c', 'N', 'N', 4096, 4096, 4096, 10);
    gemm_case('c', 'H', 'N', 4096, 4096, 4096, 10);
    gemv_case('N', 32768, 32768, 30);
    gemv_case('T', 32768, 32768, 30);
    dot_case(1L << 28, 30);
    gemm_case('d', 'N', 'N', 32768, 32768, 32768, 2);
    return 0;
}
