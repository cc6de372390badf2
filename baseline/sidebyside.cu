// baseline/sidebyside.cu -- libaccblas_b200.so and cuBLAS timed by the SAME C driver, same buffers,
// same method, one JSON line per case with both numbers.  The shapes are the reference's own
// benchmark shapes (bench/Bench/Accelerate.hs:44-58,92-125: gemm n = 100, 250, 500, 1000;
// gemv / gevm 200x400 .. 2000x3000; four element types) plus the BASELINE.json configs.
// cuBLAS is called the way the reference calls it (operand swap, beta = 0, LLVM/PTX/Level3.hs:91-106,
// LLVM/PTX/Level2.hs:105-124).  REPORTED BASELINE TOOL: nothing here is part of the product.
//
// usage: sidebyside [filter-substring | refshapes | ksweep | tiles | families] [iters]
#include <cublas_v2.h>
#include <cuda_runtime.h>
#include <algorithm>
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>
#include "../include/accblas_b200.h"

#define CK(x) do { auto e_ = (x); if (e_ != 0) { fprintf(stderr, "fail %s -> %d line %d\n", #x, (int)e_, __LINE__); exit(1); } } while (0)

__global__ void fill(double *p, size_t n, unsigned seed) {
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
static const char DT[4] = {'s', 'd', 'c', 'z'};            // accblas_dtype_t order
static size_t esz(int dt) { return dt == 0 ? 4 : dt == 3 ? 16 : 8; }

static void *dev_random(int dt, size_t elems, unsigned seed) {
    void *p; CK(cudaMalloc(&p, esz(dt) * std::max<size_t>(elems, 1)));
    if (dt == 0 || dt == 2) fillf<<<592, 256>>>((float *)p, esz(dt) / 4 * elems, seed);
    else fill<<<592, 256>>>((double *)p, esz(dt) / 8 * elems, seed);
    return p;
}

// calls enqueued back to back, one event between calls (tools/perf.py and cublas_ref.cu use the same method)
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

// sustained time per call: `iters` calls enqueued back to back between TWO events (no event between calls; the per-call
// events above cost the short kernels a visible gap and sit on the timer's ~1 us quantum).  Best of 3 batches.
static double now_ms();
static float g_enqueue_us = 0.f;                 // host time per call of the last stream_ms batch (enqueue only, no sync)
template <typename F> float stream_ms(F f, int iters) {
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    float best = 1e30f;
    g_enqueue_us = 1e30f;
    for (int rep = 0; rep < 3; ++rep) {
        CK(cudaDeviceSynchronize());
        cudaEventRecord(e0);
        const double h0 = now_ms();
        for (int i = 0; i < iters; ++i) f();
        const double h1 = now_ms();
        cudaEventRecord(e1);
        CK(cudaDeviceSynchronize());
        float t; cudaEventElapsedTime(&t, e0, e1);
        best = std::min(best, t / iters);
        g_enqueue_us = std::min(g_enqueue_us, (float)((h1 - h0) * 1e3 / iters));
    }
    cudaEventDestroy(e0); cudaEventDestroy(e1);
    return best;
}

static cublasOperation_t enc(char c) { return c == 'N' ? CUBLAS_OP_N : c == 'T' ? CUBLAS_OP_T : CUBLAS_OP_C; }
static int opcode(char c) { return c == 'N' ? 0 : c == 'T' ? 1 : 2; }

static void gemm_case(int dt, char opA, char opB, long m, long n, long k, int iters) {
    char name[80]; snprintf(name, sizeof name, "%cgemm_%c%c_%ldx%ldx%ld", DT[dt], opA, opB, m, n, k);
    if (!strstr(name, filter) && strcmp(filter, "ksweep")) return;
    if (iters_override) iters = iters_override;
    void *A = dev_random(dt, (size_t)m * k, 1), *B = dev_random(dt, (size_t)k * n, 2), *C = dev_random(dt, (size_t)m * n, 3);
    long rowsA = opA == 'N' ? m : k, colsA = opA == 'N' ? k : m, rowsB = opB == 'N' ? k : n, colsB = opB == 'N' ? n : k;
    int lda = (int)colsA, ldb = (int)colsB;
    double one[2] = {1.0, 0.0}; float onef[2] = {1.f, 0.f};
    const void *alpha = (dt == 0 || dt == 2) ? (const void *)onef : (const void *)one;
    auto ours = [&]() { CK(accblas_gemm(nullptr, dt, opcode(opA), opcode(opB), rowsA, colsA, rowsB, colsB, alpha, A, B, C)); };
    auto theirs = [&]() {
        if (dt == 0) { float al = 1, be = 0; CK(cublasSgemm(H, enc(opB), enc(opA), n, m, k, &al, (float *)B, ldb, (float *)A, lda, &be, (float *)C, n)); }
        if (dt == 1) { double al = 1, be = 0; CK(cublasDgemm(H, enc(opB), enc(opA), n, m, k, &al, (double *)B, ldb, (double *)A, lda, &be, (double *)C, n)); }
        if (dt == 2) { cuComplex al = {1, 0}, be = {0, 0}; CK(cublasCgemm(H, enc(opB), enc(opA), n, m, k, &al, (cuComplex *)B, ldb, (cuComplex *)A, lda, &be, (cuComplex *)C, n)); }
        if (dt == 3) { cuDoubleComplex al = {1, 0}, be = {0, 0}; CK(cublasZgemm(H, enc(opB), enc(opA), n, m, k, &al, (cuDoubleComplex *)B, ldb, (cuDoubleComplex *)A, lda, &be, (cuDoubleComplex *)C, n)); }
    };
    float b1, b2, m1 = time_ms(ours, 5, iters, &b1), m2 = time_ms(theirs, 5, iters, &b2);
    double flop = (dt >= 2 ? 8.0 : 2.0) * m * n * k;
    printf("{\"case\":\"%s\",\"kernel\":\"%s\",\"accblas_ms\":%.4f,\"accblas_ms_best\":%.4f,\"cublas_ms\":%.4f,\"cublas_ms_best\":%.4f,"
           "\"accblas_tflops\":%.3f,\"cublas_tflops\":%.3f,\"speedup\":%.3f",
           name, accblas_gemm_kernel_name(dt, opcode(opA), opcode(opB), rowsA, colsA, rowsB, colsB), m1, b1, m2, b2,
           flop / m1 * 1e-9, flop / m2 * 1e-9, m2 / m1);
    if (m1 < 1.0f) {
        const float s1 = stream_ms(ours, iters), q1 = g_enqueue_us, s2 = stream_ms(theirs, iters), q2 = g_enqueue_us;
        printf(",\"accblas_ms_stream\":%.4f,\"cublas_ms_stream\":%.4f,\"speedup_stream\":%.3f,\"accblas_enqueue_us\":%.2f,\"cublas_enqueue_us\":%.2f",
               s1, s2, s2 / s1, q1, q2);
    }
    printf("}\n");
    fflush(stdout);
    cudaFree(A); cudaFree(B); cudaFree(C);
}

// op: 'N' = (#>), 'T' = (<#)
static void gemv_case(int dt, char op, long rows, long cols, int iters) {
    char name[80]; snprintf(name, sizeof name, "%cgemv_%c_%ldx%ld", DT[dt], op, rows, cols);
    if (!strstr(name, filter)) return;
    if (iters_override) iters = iters_override;
    long lenx = op == 'N' ? cols : rows, leny = op == 'N' ? rows : cols;
    void *A = dev_random(dt, (size_t)rows * cols, 3), *x = dev_random(dt, lenx, 4), *y = dev_random(dt, leny, 5);
    double one[2] = {1.0, 0.0}; float onef[2] = {1.f, 0.f};
    const void *alpha = (dt == 0 || dt == 2) ? (const void *)onef : (const void *)one;
    cublasOperation_t o = op == 'N' ? CUBLAS_OP_T : CUBLAS_OP_N;      // LLVM/PTX/Level2.hs:84-87
    auto ours = [&]() { CK(accblas_gemv(nullptr, dt, opcode(op), rows, cols, alpha, A, x, lenx, y)); };
    auto theirs = [&]() {
        if (dt == 0) { float al = 1, be = 0; CK(cublasSgemv(H, o, (int)cols, (int)rows, &al, (float *)A, (int)cols, (float *)x, 1, &be, (float *)y, 1)); }
        if (dt == 1) { double al = 1, be = 0; CK(cublasDgemv(H, o, (int)cols, (int)rows, &al, (double *)A, (int)cols, (double *)x, 1, &be, (double *)y, 1)); }
        if (dt == 2) { cuComplex al = {1, 0}, be = {0, 0}; CK(cublasCgemv(H, o, (int)cols, (int)rows, &al, (cuComplex *)A, (int)cols, (cuComplex *)x, 1, &be, (cuComplex *)y, 1)); }
        if (dt == 3) { cuDoubleComplex al = {1, 0}, be = {0, 0}; CK(cublasZgemv(H, o, (int)cols, (int)rows, &al, (cuDoubleComplex *)A, (int)cols, (cuDoubleComplex *)x, 1, &be, (cuDoubleComplex *)y, 1)); }
    };
    float b1, b2, m1 = time_ms(ours, 5, iters, &b1), m2 = time_ms(theirs, 5, iters, &b2);
    double bytes = (double)esz(dt) * ((double)rows * cols + lenx + leny);
    printf("{\"case\":\"%s\",\"accblas_ms\":%.4f,\"accblas_ms_best\":%.4f,\"cublas_ms\":%.4f,\"cublas_ms_best\":%.4f,"
           "\"accblas_gbs\":%.1f,\"cublas_gbs\":%.1f,\"speedup\":%.3f",
           name, m1, b1, m2, b2, bytes / m1 * 1e-6, bytes / m2 * 1e-6, m2 / m1);
    if (m1 < 1.0f) {
        const float s1 = stream_ms(ours, iters), q1 = g_enqueue_us, s2 = stream_ms(theirs, iters), q2 = g_enqueue_us;
        printf(",\"accblas_ms_stream\":%.4f,\"cublas_ms_stream\":%.4f,\"speedup_stream\":%.3f,\"accblas_enqueue_us\":%.2f,\"cublas_enqueue_us\":%.2f",
               s1, s2, s2 / s1, q1, q2);
    }
    printf("}\n");
    fflush(stdout);
    cudaFree(A); cudaFree(x); cudaFree(y);
}

static void dot_case(int dt, long n, int iters) {
    char name[80]; snprintf(name, sizeof name, "%cdot_%ld", DT[dt], n);
    if (!strstr(name, filter)) return;
    if (iters_override) iters = iters_override;
    void *x = dev_random(dt, n, 5), *y = dev_random(dt, n, 6), *out; CK(cudaMalloc(&out, 16));
    CK(cublasSetPointerMode(H, CUBLAS_POINTER_MODE_DEVICE));
    auto ours = [&]() { CK(accblas_dot(nullptr, dt, 0, n, x, y, out)); };
    auto theirs = [&]() {
        if (dt == 0) CK(cublasSdot(H, (int)n, (float *)x, 1, (float *)y, 1, (float *)out));
        if (dt == 1) CK(cublasDdot(H, (int)n, (double *)x, 1, (double *)y, 1, (double *)out));
        if (dt == 2) CK(cublasCdotu(H, (int)n, (cuComplex *)x, 1, (cuComplex *)y, 1, (cuComplex *)out));
        if (dt == 3) CK(cublasZdotu(H, (int)n, (cuDoubleComplex *)x, 1, (cuDoubleComplex *)y, 1, (cuDoubleComplex *)out));
    };
    float b1, b2, m1 = time_ms(ours, 5, iters, &b1), m2 = time_ms(theirs, 5, iters, &b2);
    const float s1 = stream_ms(ours, iters), s2 = stream_ms(theirs, iters);
    CK(cublasSetPointerMode(H, CUBLAS_POINTER_MODE_HOST));
    double bytes = 2.0 * esz(dt) * n + esz(dt);
    printf("{\"case\":\"%s\",\"note\":\"cuBLAS dot is a stand-in: the reference's dot is Accelerate's fold\",\"accblas_ms\":%.4f,\"accblas_ms_best\":%.4f,"
           "\"cublas_ms\":%.4f,\"cublas_ms_best\":%.4f,\"accblas_gbs\":%.1f,\"cublas_gbs\":%.1f,\"speedup\":%.3f,"
           "\"accblas_ms_stream\":%.4f,\"cublas_ms_stream\":%.4f,\"speedup_stream\":%.3f}\n",
           name, m1, b1, m2, b2, bytes / m1 * 1e-6, bytes / m2 * 1e-6, m2 / m1, s1, s2, s2 / s1);
    fflush(stdout);
    cudaFree(x); cudaFree(y); cudaFree(out);
}

static double now_ms() {
    return std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now().time_since_epoch()).count();
}

// End to end from pinned host buffers: accblas_gemm_host (wavefront pipeline) against the reference's
// `run` flow restated with cuBLAS: upload A, upload B, cublas?gemm, download C, one after the other.
static void host_case(int dt, char opA, char opB, long m, long n, long k, int iters) {
    char name[80]; snprintf(name, sizeof name, "host_%cgemm_%c%c_%ldx%ldx%ld", DT[dt], opA, opB, m, n, k);
    if (!strstr(name, filter)) return;
    if (iters_override) iters = iters_override;
    const size_t es = esz(dt), bA = es * m * k, bB = es * k * n, bC = es * m * n;
    void *dA = dev_random(dt, (size_t)m * k, 1), *dB = dev_random(dt, (size_t)k * n, 2), *dC; CK(cudaMalloc(&dC, bC));
    void *hA, *hB, *hC; CK(cudaMallocHost(&hA, bA)); CK(cudaMallocHost(&hB, bB)); CK(cudaMallocHost(&hC, bC));
    CK(cudaMemcpy(hA, dA, bA, cudaMemcpyDeviceToHost)); CK(cudaMemcpy(hB, dB, bB, cudaMemcpyDeviceToHost));
    long rowsA = opA == 'N' ? m : k, colsA = opA == 'N' ? k : m, rowsB = opB == 'N' ? k : n, colsB = opB == 'N' ? n : k;
    double one[2] = {1.0, 0.0}; float onef[2] = {1.f, 0.f};
    const void *alpha = (dt == 0 || dt == 2) ? (const void *)onef : (const void *)one;
    // raw link rates
    double t0 = now_ms(); CK(cudaMemcpy(dA, hA, bA, cudaMemcpyHostToDevice)); double h2d = bA / (now_ms() - t0) * 1e-6;
    t0 = now_ms(); CK(cudaMemcpy(hC, dC, bC, cudaMemcpyDeviceToHost)); double d2h = bC / (now_ms() - t0) * 1e-6;
    std::vector<double> ours, theirs;
    for (int i = 0; i < iters + 2; ++i) {
        t0 = now_ms();
        CK(accblas_gemm_host(nullptr, dt, opcode(opA), opcode(opB), rowsA, colsA, rowsB, colsB, alpha, hA, hB, hC));
        if (i >= 2) ours.push_back(now_ms() - t0);
    }
    int lda = (int)colsA, ldb = (int)colsB;
    for (int i = 0; i < iters + 2; ++i) {
        t0 = now_ms();
        CK(cudaMemcpyAsync(dA, hA, bA, cudaMemcpyHostToDevice, 0));
        CK(cudaMemcpyAsync(dB, hB, bB, cudaMemcpyHostToDevice, 0));
        if (dt == 0) { float al = 1, be = 0; CK(cublasSgemm(H, enc(opB), enc(opA), n, m, k, &al, (float *)dB, ldb, (float *)dA, lda, &be, (float *)dC, n)); }
        if (dt == 1) { double al = 1, be = 0; CK(cublasDgemm(H, enc(opB), enc(opA), n, m, k, &al, (double *)dB, ldb, (double *)dA, lda, &be, (double *)dC, n)); }
        if (dt == 2) { cuComplex al = {1, 0}, be = {0, 0}; CK(cublasCgemm(H, enc(opB), enc(opA), n, m, k, &al, (cuComplex *)dB, ldb, (cuComplex *)dA, lda, &be, (cuComplex *)dC, n)); }
        if (dt == 3) { cuDoubleComplex al = {1, 0}, be = {0, 0}; CK(cublasZgemm(H, enc(opB), enc(opA), n, m, k, &al, (cuDoubleComplex *)dB, ldb, (cuDoubleComplex *)dA, lda, &be, (cuDoubleComplex *)dC, n)); }
        CK(cudaMemcpyAsync(hC, dC, bC, cudaMemcpyDeviceToHost, 0));
        CK(cudaStreamSynchronize(0));
        if (i >= 2) theirs.push_back(now_ms() - t0);
    }
    std::sort(ours.begin(), ours.end()); std::sort(theirs.begin(), theirs.end());
    double m1 = ours[ours.size() / 2], m2 = theirs[theirs.size() / 2];
    double flop = (dt >= 2 ? 8.0 : 2.0) * m * n * k;
    printf("{\"case\":\"%s\",\"h2d_gbs\":%.1f,\"d2h_gbs\":%.1f,\"accblas_host_ms\":%.3f,\"accblas_host_ms_best\":%.3f,"
           "\"cublas_serial_ms\":%.3f,\"accblas_tflops\":%.3f,\"cublas_serial_tflops\":%.3f,\"speedup\":%.3f}\n",
           name, h2d, d2h, m1, ours[0], m2, flop / m1 * 1e-9, flop / m2 * 1e-9, m2 / m1);
    fflush(stdout);
    cudaFree(dA); cudaFree(dB); cudaFree(dC); cudaFreeHost(hA); cudaFreeHost(hB); cudaFreeHost(hC);
}

// every CTA tile of the DMMA family on a ladder of sizes: calibrates the tile-choice cost model
static void tile_sweep() {
    for (int dt : {1, 3}) {
        const int knob = dt == 1 ? ACCBLAS_TUNE_DMMA_TILE : ACCBLAS_TUNE_ZMMA_TILE, ntiles = dt == 1 ? 4 : 3;
        for (long n : {100L, 160L, 250L, 384L, 500L, 512L, 768L, 1000L, 1024L, 1280L, 1536L, 2048L, 3072L, 4096L, 8192L}) {
            if (dt == 3 && n > 4096) continue;
            if (dt == 1 && n == 8192 && 0) continue;
            void *A = dev_random(dt, (size_t)n * n, 1), *B = dev_random(dt, (size_t)n * n, 2), *C = dev_random(dt, (size_t)n * n, 3);
            double one[2] = {1.0, 0.0};
            printf("{\"sweep\":\"%cgemm_NN\",\"n\":%ld", DT[dt], n);
            for (int tile = -1; tile < ntiles; ++tile) {
                accblas_set_tuning(knob, tile);
                auto ours = [&]() { CK(accblas_gemm(nullptr, dt, 0, 0, n, n, n, n, one, A, B, C)); };
                float best, med = time_ms(ours, 3, n <= 1024 ? 100 : 20, &best);
                if (tile < 0) printf(",\"auto_us\":%.1f", med * 1e3); else printf(",\"tile%d_us\":%.1f", tile, med * 1e3);
            }
            accblas_set_tuning(knob, -1);
            auto theirs = [&]() {
                if (dt == 1) { double al = 1, be = 0; CK(cublasDgemm(H, CUBLAS_OP_N, CUBLAS_OP_N, n, n, n, &al, (double *)B, n, (double *)A, n, &be, (double *)C, n)); }
                else { cuDoubleComplex al = {1, 0}, be = {0, 0}; CK(cublasZgemm(H, CUBLAS_OP_N, CUBLAS_OP_N, n, n, n, &al, (cuDoubleComplex *)B, n, (cuDoubleComplex *)A, n, &be, (cuDoubleComplex *)C, n)); }
            };
            float best, med = time_ms(theirs, 3, n <= 1024 ? 100 : 20, &best);
            printf(",\"cublas_us\":%.1f}\n", med * 1e3);
            fflush(stdout);
            cudaFree(A); cudaFree(B); cudaFree(C);
        }
    }
}

// f32 / c32: mma.sync family (each tile) against the tcgen05 family and cuBLAS on a ladder of sizes
static void family_sweep() {
    for (int dt : {0, 2}) {
        for (long n : {100L, 160L, 250L, 384L, 500L, 768L, 1000L, 1024L, 1536L, 2048L, 3072L, 4096L}) {
            if (dt == 2 && n > 3072) continue;
            void *A = dev_random(dt, (size_t)n * n, 1), *B = dev_random(dt, (size_t)n * n, 2), *C = dev_random(dt, (size_t)n * n, 3);
            float one[2] = {1.f, 0.f};
            auto ours = [&]() { CK(accblas_gemm(nullptr, dt, 0, 0, n, n, n, n, one, A, B, C)); };
            const int it = n <= 1024 ? 100 : 20;
            float best;
            printf("{\"sweep\":\"%cgemm_NN\",\"n\":%ld", DT[dt], n);
            printf(",\"auto_us\":%.1f,\"auto_kernel\":\"%s\"", time_ms(ours, 3, it, &best) * 1e3, accblas_gemm_kernel_name(dt, 0, 0, n, n, n, n));
            accblas_set_tuning(ACCBLAS_TUNE_SGEMM_FAMILY, 1);
            for (int tile = 0; tile < 3; ++tile) {
                if (dt == 2 && tile == 1) continue;
                accblas_set_tuning(ACCBLAS_TUNE_SMMA_TILE, tile);
                printf(",\"mma_tile%d_us\":%.1f", tile, time_ms(ours, 3, it, &best) * 1e3);
            }
            accblas_set_tuning(ACCBLAS_TUNE_SMMA_TILE, -1);
            accblas_set_tuning(ACCBLAS_TUNE_SGEMM_FAMILY, 2);
            if (strstr(accblas_gemm_kernel_name(dt, 0, 0, n, n, n, n), "tcgen05"))
                for (int tile = 0; tile < 3; ++tile) {               // 128 x 256 / 128 x 128 / 128 x 64
                    accblas_set_tuning(ACCBLAS_TUNE_TC_TILE, tile);
                    printf(",\"tcgen05_tile%d_us\":%.1f", tile, time_ms(ours, 3, it, &best) * 1e3);
                }
            accblas_set_tuning(ACCBLAS_TUNE_TC_TILE, -1);
            accblas_set_tuning(ACCBLAS_TUNE_SGEMM_FAMILY, -1);
            auto theirs = [&]() {
                if (dt == 0) { float al = 1, be = 0; CK(cublasSgemm(H, CUBLAS_OP_N, CUBLAS_OP_N, n, n, n, &al, (float *)B, n, (float *)A, n, &be, (float *)C, n)); }
                else { cuComplex al = {1, 0}, be = {0, 0}; CK(cublasCgemm(H, CUBLAS_OP_N, CUBLAS_OP_N, n, n, n, &al, (cuComplex *)B, n, (cuComplex *)A, n, &be, (cuComplex *)C, n)); }
            };
            printf(",\"cublas_us\":%.1f}\n", time_ms(theirs, 3, it, &best) * 1e3);
            fflush(stdout);
            cudaFree(A); cudaFree(B); cudaFree(C);
        }
    }
}

int main(int argc, char **argv) {
    if (argc > 1) filter = argv[1];
    if (argc > 2) iters_override = atoi(argv[2]);
    CK(cublasCreate(&H));
    CK(cublasSetAtomicsMode(H, CUBLAS_ATOMICS_ALLOWED));              // LLVM/PTX/Context.hs:59
    if (!strcmp(filter, "ksweep")) {                                  // fixed m = n, growing k: fixed cost vs per-K cost of the small kernels
        for (int dt = 0; dt < 4; ++dt)
            for (long mn : {96L, 250L})
                for (long k : {16L, 64L, 128L, 256L, 512L, 1024L}) gemm_case(dt, 'N', 'N', mn, mn, k, 300);
        return 0;
    }
    if (!strcmp(filter, "k1024")) { filter = ""; gemm_case(1, 'N', 'N', 250, 250, 1024, 20); gemm_case(1, 'N', 'N', 250, 250, 250, 20); gemm_case(1, 'N', 'N', 100, 100, 100, 20); return 0; }
    if (!strcmp(filter, "tiles")) { tile_sweep(); return 0; }
    if (!strcmp(filter, "families")) { family_sweep(); return 0; }
    // "refshapes": only the 16 gemm + 32 gemv / gevm cases the reference itself benchmarks (bench.py's extras run this)
    const bool ref_only = !strcmp(filter, "refshapes");
    if (ref_only) filter = "";
    // --- the reference's own benchmark shapes (bench/Bench/Accelerate.hs) ---
    for (int dt = 0; dt < 4; ++dt)
        for (long n : {100L, 250L, 500L, 1000L}) gemm_case(dt, 'N', 'N', n, n, n, 200);
    for (int dt = 0; dt < 4; ++dt)
        for (auto sh : {std::pair<long, long>{200, 400}, {500, 1000}, {1000, 2000}, {2000, 3000}}) {
            gemv_case(dt, 'N', sh.first, sh.second, 200);
            gemv_case(dt, 'T', sh.first, sh.second, 200);
        }
    if (ref_only) return 0;
    // --- BASELINE.json configs and the sizes in between ---
    for (long n : {512L, 1024L, 2048L, 4096L}) gemm_case(1, 'N', 'N', n, n, n, 50);
    for (long n : {2048L, 4096L}) gemm_case(0, 'N', 'N', n, n, n, 50);
    for (const char *ops : {"NN", "NT", "TN"}) {
        gemm_case(1, ops[0], ops[1], 8192, 8192, 8192, 10);
        gemm_case(0, ops[0], ops[1], 8192, 8192, 8192, 10);
    }
    gemm_case(3, 'N', 'N', 4096, 4096, 4096, 10);
    gemm_case(3, 'H', 'N', 4096, 4096, 4096, 10);
    gemm_case(2, 'H', 'N', 4096, 4096, 4096, 10);
    gemv_case(1, 'N', 32768, 32768, 30);
    gemv_case(1, 'T', 32768, 32768, 30);
    dot_case(1, 1L << 28, 30);
    for (int dt = 0; dt < 4; ++dt) dot_case(dt, 1L << 20, 200);
    // --- end to end from pinned host memory ---
    host_case(1, 'N', 'N', 8192, 8192, 8192, 5);
    host_case(1, 'T', 'N', 8192, 8192, 8192, 3);
    host_case(0, 'N', 'N', 8192, 8192, 8192, 5);
    host_case(1, 'N', 'N', 1000, 1000, 1000, 20);
    return 0;
}
