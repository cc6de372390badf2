// tools/exp/gemv_n_exp.cu -- design-space probe for the row-major gemv N kernel (f64), standalone.
// Not part of the library: variants are timed against each other on one B200 and the winner is
// moved into csrc/gemv.cu.   nvcc -O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -o gemv_n_exp gemv_n_exp.cu
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdint>
#include <cstdlib>
#include <cmath>
#include <vector>
#include <algorithm>

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { printf("CUDA %s at %s:%d\n", cudaGetErrorString(e_), __FILE__, __LINE__); exit(1); } } while (0)

__device__ __forceinline__ int4 ldg_stream(const void *p) {
    int4 r;
    asm volatile("ld.global.nc.L1::no_allocate.v4.s32 {%0,%1,%2,%3}, [%4];" : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p));
    return r;
}
union V16 { int4 raw; double e[2]; };
__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
    for (int m = 16; m > 0; m >>= 1) v += __shfl_xor_sync(0xffffffffu, v, m);
    return v;
}

__global__ void fill(double *p, size_t n, uint64_t seed) {
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
        uint64_t z = (i + seed) * 0x9E3779B97F4A7C15ull; z ^= z >> 31; z *= 0xBF58476D1CE4E5B9ull; z ^= z >> 29;
        p[i] = (double)(int64_t)(z >> 11) * (1.0 / 4503599627370496.0) - 1.0;
    }
}
__global__ void ref_kernel(int64_t m, int64_t n, const double *A, const double *x, double *y) {
    int64_t row = blockIdx.x * (int64_t)(blockDim.x >> 5) + (threadIdx.x >> 5);
    if (row >= m) return;
    double s = 0;
    for (int64_t j = threadIdx.x & 31; j < n; j += 32) s = fma(A[row * n + j], x[j], s);
    s = warp_sum(s);
    if ((threadIdx.x & 31) == 0) y[row] = s;
}

// ---------------------------------------------------------------- V0: current library kernel
template <int ROWS, int THREADS, int MINB>
__global__ void __launch_bounds__(THREADS, MINB) v0(int64_t m, int64_t n, const double *__restrict__ A, const double *__restrict__ x, double *__restrict__ y) {
    __shared__ double red[ROWS][THREADS / 32];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int64_t nunits = m / ROWS, nv = n / 2;
    for (int64_t u = blockIdx.x; u < nunits; u += gridDim.x) {
        const double *row0 = A + u * ROWS * n;
        double acc[ROWS];
#pragma unroll
        for (int r = 0; r < ROWS; ++r) acc[r] = 0;
        for (int64_t v = threadIdx.x; v < nv; v += THREADS) {
            V16 xv, av[ROWS];
#pragma unroll
            for (int r = 0; r < ROWS; ++r) av[r].raw = ldg_stream(reinterpret_cast<const int4 *>(row0 + r * n) + v);
            xv.raw = __ldg(reinterpret_cast<const int4 *>(x) + v);
#pragma unroll
            for (int r = 0; r < ROWS; ++r) { acc[r] = fma(av[r].e[0], xv.e[0], acc[r]); acc[r] = fma(av[r].e[1], xv.e[1], acc[r]); }
        }
#pragma unroll
        for (int r = 0; r < ROWS; ++r) { acc[r] = warp_sum(acc[r]); if (lane == 0) red[r][warp] = acc[r]; }
        __syncthreads();
        if (threadIdx.x < ROWS) { double s = 0; for (int w = 0; w < THREADS / 32; ++w) s += red[threadIdx.x][w]; y[u * ROWS + threadIdx.x] = s; }
        __syncthreads();
    }
}

// ---------------------------------------------------------------- V1: a warp owns ROWS rows, no CTA sync
template <int ROWS, int U, int THREADS, int MINB>
__global__ void __launch_bounds__(THREADS, MINB) v1(int64_t m, int64_t n, const double *__restrict__ A, const double *__restrict__ x, double *__restrict__ y) {
    const int lane = threadIdx.x & 31;
    const int64_t gw = blockIdx.x * (int64_t)(THREADS / 32) + (threadIdx.x >> 5), nw = gridDim.x * (int64_t)(THREADS / 32);
    const int64_t nunits = m / ROWS, nv = n / 2;
    for (int64_t u = gw; u < nunits; u += nw) {
        const double *row0 = A + u * ROWS * n;
        double acc[ROWS];
#pragma unroll
        for (int r = 0; r < ROWS; ++r) acc[r] = 0;
        for (int64_t v = lane; v < nv; v += 32 * U) {
            V16 xv[U], av[ROWS][U];
#pragma unroll
            for (int r = 0; r < ROWS; ++r)
#pragma unroll
                for (int q = 0; q < U; ++q) av[r][q].raw = ldg_stream(reinterpret_cast<const int4 *>(row0 + r * n) + v + 32 * q);
#pragma unroll
            for (int q = 0; q < U; ++q) xv[q].raw = __ldg(reinterpret_cast<const int4 *>(x) + v + 32 * q);
#pragma unroll
            for (int r = 0; r < ROWS; ++r)
#pragma unroll
                for (int q = 0; q < U; ++q) { acc[r] = fma(av[r][q].e[0], xv[q].e[0], acc[r]); acc[r] = fma(av[r][q].e[1], xv[q].e[1], acc[r]); }
        }
#pragma unroll
        for (int r = 0; r < ROWS; ++r) { acc[r] = warp_sum(acc[r]); if (lane == 0) y[u * ROWS + r] = acc[r]; }
    }
}

// ---------------------------------------------------------------- V3: bulk-copy (TMA 1-D) ring, producer warp + consumer warps
__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) { asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count)); }
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes) { asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory"); }
__device__ __forceinline__ void mbar_arrive(uint32_t bar) { asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory"); }
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
    uint32_t ok = 0, spins = 0;
    while (!ok) {
        asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}" : "=r"(ok) : "r"(bar), "r"(parity) : "memory");
        if (!ok && ++spins > (1u << 26)) __trap();
    }
}
__device__ __forceinline__ void bulk_g2s(uint32_t dst, const void *src, uint32_t bytes, uint32_t bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(dst), "l"(src), "r"(bytes), "r"(bar) : "memory");
}

// CTA b owns the contiguous rows [m*b/G, m*(b+1)/G): perfectly balanced bytes.  Rows are walked in units of
// ROWS; a stage holds ROWS x CH bytes of A plus the matching CH bytes of x.
template <int ROWS, int CH, int STAGES, int CWARPS>
__global__ void __launch_bounds__(CWARPS * 32 + 32, 1) v3(int64_t m, int64_t n, const double *__restrict__ A, const double *__restrict__ x, double *__restrict__ y) {
    extern __shared__ __align__(128) unsigned char smem[];
    constexpr int STAGE_BYTES = (ROWS + 1) * CH;
    constexpr int CT = CWARPS * 32;
    constexpr int VPT = CH / 16 / CT;          // 16-byte vectors per thread per row per stage
    static_assert(VPT >= 1 && VPT * CT * 16 == CH, "CH must be a multiple of 16 * consumer threads");
    uint64_t *full = reinterpret_cast<uint64_t *>(smem + STAGES * STAGE_BYTES);
    uint64_t *empty = full + STAGES;
    double *red = reinterpret_cast<double *>(empty + STAGES);   // [ROWS][CWARPS]
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (threadIdx.x == 0) {
        for (int s = 0; s < STAGES; ++s) { mbar_init(smem_u32(full + s), 1); mbar_init(smem_u32(empty + s), CWARPS); }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    const int64_t r_begin = m * blockIdx.x / gridDim.x, r_end = m * (blockIdx.x + 1) / gridDim.x;
    const int64_t row_bytes = n * 8;
    const int64_t chunks = row_bytes / CH;     // host guarantees divisibility
    if (warp == CWARPS) {
        if (lane == 0) {
            uint32_t it = 0;
            for (int64_t r0 = r_begin; r0 < r_end; r0 += ROWS) {
                const int valid = (int)((r_end - r0) < ROWS ? (r_end - r0) : ROWS);
                for (int64_t c = 0; c < chunks; ++c, ++it) {
                    const int s = it % STAGES;
                    const uint32_t ph = (it / STAGES) & 1;
                    mbar_wait(smem_u32(empty + s), ph ^ 1);
                    const uint32_t fb = smem_u32(full + s);
                    mbar_expect_tx(fb, (uint32_t)(valid + 1) * CH);
                    const uint32_t base = smem_u32(smem + s * STAGE_BYTES);
                    const char *src = reinterpret_cast<const char *>(A) + r0 * row_bytes + c * CH;
                    for (int r = 0; r < valid; ++r) bulk_g2s(base + r * CH, src + r * row_bytes, CH, fb);
                    bulk_g2s(base + ROWS * CH, reinterpret_cast<const char *>(x) + c * CH, CH, fb);
                }
            }
        }
        return;
    }
    uint32_t it = 0;
    for (int64_t r0 = r_begin; r0 < r_end; r0 += ROWS) {
        double acc[ROWS];
#pragma unroll
        for (int r = 0; r < ROWS; ++r) acc[r] = 0;
        for (int64_t c = 0; c < chunks; ++c, ++it) {
            const int s = it % STAGES;
            const uint32_t ph = (it / STAGES) & 1;
            mbar_wait(smem_u32(full + s), ph);
            const unsigned char *st = smem + s * STAGE_BYTES;
#pragma unroll
            for (int q = 0; q < VPT; ++q) {
                const int off = (q * CT + threadIdx.x) * 16;
                V16 xv; xv.raw = *reinterpret_cast<const int4 *>(st + ROWS * CH + off);
#pragma unroll
                for (int r = 0; r < ROWS; ++r) {
                    V16 av; av.raw = *reinterpret_cast<const int4 *>(st + r * CH + off);
                    acc[r] = fma(av.e[0], xv.e[0], acc[r]);
                    acc[r] = fma(av.e[1], xv.e[1], acc[r]);
                }
            }
            __syncwarp();
            if (lane == 0) mbar_arrive(smem_u32(empty + s));
        }
#pragma unroll
        for (int r = 0; r < ROWS; ++r) { acc[r] = warp_sum(acc[r]); if (lane == 0) red[r * CWARPS + warp] = acc[r]; }
        asm volatile("bar.sync 1, %0;" ::"n"(CT));
        if (threadIdx.x < ROWS && r0 + threadIdx.x < r_end) {
            double sum = 0;
#pragma unroll
            for (int w = 0; w < CWARPS; ++w) sum += red[threadIdx.x * CWARPS + w];
            y[r0 + threadIdx.x] = sum;
        }
        asm volatile("bar.sync 1, %0;" ::"n"(CT));
    }
}

// ---------------------------------------------------------------- V4: V0 with balanced contiguous row ranges and no register cap
template <int ROWS, int THREADS, int MINB>
__global__ void __launch_bounds__(THREADS, MINB) v4(int64_t m, int64_t n, const double *__restrict__ A, const double *__restrict__ x, double *__restrict__ y) {
    __shared__ double red[ROWS][THREADS / 32];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int64_t r_begin = m * blockIdx.x / gridDim.x, r_end = m * (blockIdx.x + 1) / gridDim.x, nv = n / 2;
    for (int64_t r0 = r_begin; r0 < r_end; r0 += ROWS) {
        const double *row0 = A + r0 * n;
        int64_t roff[ROWS];
#pragma unroll
        for (int r = 0; r < ROWS; ++r) roff[r] = (r0 + r < r_end ? r : 0) * n;    // tail rows re-read row r0 (L2 hit)
        double acc[ROWS];
#pragma unroll
        for (int r = 0; r < ROWS; ++r) acc[r] = 0;
        for (int64_t v = threadIdx.x; v < nv; v += THREADS) {
            V16 xv, av[ROWS];
#pragma unroll
            for (int r = 0; r < ROWS; ++r) av[r].raw = ldg_stream(reinterpret_cast<const int4 *>(row0 + roff[r]) + v);
            xv.raw = __ldg(reinterpret_cast<const int4 *>(x) + v);
#pragma unroll
            for (int r = 0; r < ROWS; ++r) { acc[r] = fma(av[r].e[0], xv.e[0], acc[r]); acc[r] = fma(av[r].e[1], xv.e[1], acc[r]); }
        }
#pragma unroll
        for (int r = 0; r < ROWS; ++r) { acc[r] = warp_sum(acc[r]); if (lane == 0) red[r][warp] = acc[r]; }
        __syncthreads();
        if (threadIdx.x < ROWS && r0 + threadIdx.x < r_end) { double s = 0; for (int w = 0; w < THREADS / 32; ++w) s += red[threadIdx.x][w]; y[r0 + threadIdx.x] = s; }
        __syncthreads();
    }
}

struct Ctx { int64_t m, n; double *A, *x, *y, *yref; std::vector<double> href; };

template <typename F> void run(const char *name, Ctx &c, F launch) {
    CK(cudaMemset(c.y, 0xff, c.m * 8));
    launch();
    CK(cudaDeviceSynchronize());
    std::vector<double> h(c.m);
    CK(cudaMemcpy(h.data(), c.y, c.m * 8, cudaMemcpyDeviceToHost));
    double worst = 0;
    for (int64_t i = 0; i < c.m; ++i) { double d = fabs(h[i] - c.href[i]); if (!(d <= worst)) worst = d; }
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    for (int i = 0; i < 3; ++i) launch();
    std::vector<float> ts;
    for (int i = 0; i < 20; ++i) { cudaEventRecord(e0); launch(); cudaEventRecord(e1); cudaEventSynchronize(e1); float ms; cudaEventElapsedTime(&ms, e0, e1); ts.push_back(ms); }
    CK(cudaGetLastError());
    std::sort(ts.begin(), ts.end());
    const double bytes = 8.0 * ((double)c.m * c.n + c.m + c.n);
    printf("{\"variant\":\"%s\",\"m\":%lld,\"n\":%lld,\"ms_median\":%.4f,\"ms_best\":%.4f,\"gbs_median\":%.1f,\"gbs_best\":%.1f,\"max_abs_err\":%.3e}\n",
           name, (long long)c.m, (long long)c.n, ts[10], ts[0], bytes / ts[10] * 1e-6, bytes / ts[0] * 1e-6, worst);
    fflush(stdout);
}

template <int ROWS, int CH, int STAGES, int CWARPS> void run_v3(const char *name, Ctx &c, int ctas_per_sm) {
    const size_t smem = (size_t)STAGES * (ROWS + 1) * CH + 2 * STAGES * 8 + ROWS * CWARPS * 8 + 64;
    CK(cudaFuncSetAttribute(v3<ROWS, CH, STAGES, CWARPS>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    run(name, c, [&] { v3<ROWS, CH, STAGES, CWARPS><<<148 * ctas_per_sm, CWARPS * 32 + 32, smem>>>(c.m, c.n, c.A, c.x, c.y); });
}

int main(int argc, char **argv) {
    Ctx c; c.m = argc > 1 ? atoll(argv[1]) : 32768; c.n = argc > 2 ? atoll(argv[2]) : 32768;
    CK(cudaMalloc(&c.A, c.m * c.n * 8)); CK(cudaMalloc(&c.x, c.n * 8)); CK(cudaMalloc(&c.y, c.m * 8)); CK(cudaMalloc(&c.yref, c.m * 8));
    fill<<<148 * 8, 256>>>(c.A, (size_t)c.m * c.n, 1); fill<<<64, 256>>>(c.x, c.n, 99);
    ref_kernel<<<(unsigned)((c.m + 7) / 8), 256>>>(c.m, c.n, c.A, c.x, c.yref);
    CK(cudaDeviceSynchronize());
    c.href.resize(c.m); CK(cudaMemcpy(c.href.data(), c.yref, c.m * 8, cudaMemcpyDeviceToHost));

    run("v0_r8_t512_b2", c, [&] { v0<8, 512, 2><<<296, 512>>>(c.m, c.n, c.A, c.x, c.y); });
    run("v0_r8_t256_b3", c, [&] { v0<8, 256, 3><<<148 * 3, 256>>>(c.m, c.n, c.A, c.x, c.y); });
    run("v0_r8_t512_b1", c, [&] { v0<8, 512, 1><<<148, 512>>>(c.m, c.n, c.A, c.x, c.y); });
    run("v0_r4_t512_b2", c, [&] { v0<4, 512, 2><<<296, 512>>>(c.m, c.n, c.A, c.x, c.y); });
    run("v0_r16_t256_b2", c, [&] { v0<16, 256, 2><<<296, 256>>>(c.m, c.n, c.A, c.x, c.y); });
    run("v4_r8_t512_b1", c, [&] { v4<8, 512, 1><<<148, 512>>>(c.m, c.n, c.A, c.x, c.y); });
    run("v4_r8_t256_b3", c, [&] { v4<8, 256, 3><<<148 * 3, 256>>>(c.m, c.n, c.A, c.x, c.y); });
    run("v4_r8_t256_b2", c, [&] { v4<8, 256, 2><<<148 * 2, 256>>>(c.m, c.n, c.A, c.x, c.y); });
    run("v4_r4_t256_b4", c, [&] { v4<4, 256, 4><<<148 * 4, 256>>>(c.m, c.n, c.A, c.x, c.y); });
    run("v1_r4_u2_t256_b3", c, [&] { v1<4, 2, 256, 3><<<148 * 3, 256>>>(c.m, c.n, c.A, c.x, c.y); });
    run("v1_r8_u1_t256_b3", c, [&] { v1<8, 1, 256, 3><<<148 * 3, 256>>>(c.m, c.n, c.A, c.x, c.y); });
    run("v1_r2_u4_t256_b4", c, [&] { v1<2, 4, 256, 4><<<148 * 4, 256>>>(c.m, c.n, c.A, c.x, c.y); });
    run_v3<8, 4096, 5, 4>("v3_r8_ch4096_s5_w4_x1", c, 1);
    run_v3<8, 2048, 5, 4>("v3_r8_ch2048_s5_w4_x2", c, 2);
    run_v3<8, 2048, 10, 4>("v3_r8_ch2048_s10_w4_x1", c, 1);
    run_v3<16, 2048, 6, 4>("v3_r16_ch2048_s6_w4_x1", c, 1);
    run_v3<16, 2048, 3, 4>("v3_r16_ch2048_s3_w4_x2", c, 2);
    run_v3<8, 8192, 3, 8>("v3_r8_ch8192_s3_w8_x1", c, 1);
    run_v3<4, 4096, 5, 4>("v3_r4_ch4096_s5_w4_x2", c, 2);
    return 0;
}
