// tools/exp/launch_floor.cu -- what a back-to-back launch costs on this GPU as a function of the kernel's shape:
// sustained time per launch (N launches between two events) of kernels that do (almost) nothing.
//   nvcc -O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -I accelerate-blas_b200/csrc -o /tmp/launch_floor tools/exp/launch_floor.cu
#include <cuda.h>
#include <cuda_runtime.h>
#include <cstdio>
#include <algorithm>
#include "tma.cuh"
using namespace accblas;

__global__ void k_empty(int *p) { if (p && threadIdx.x == 9999) *p = 1; }
__global__ void k_dynsmem(int *p) { extern __shared__ uint8_t s[]; if (p && threadIdx.x == 9999) *p = s[0]; }
struct Big { char pad[384]; };
__global__ void k_bigparam(const __grid_constant__ Big b, int *p) { if (p && threadIdx.x == 9999) *p = b.pad[3]; }
// first data through TMA: barrier init, descriptor prefetch, one 16 x 128-byte box, wait
__global__ void k_tma_first(const __grid_constant__ CUtensorMap map, double *out, int boxes)
{
    extern __shared__ uint8_t smem_raw[];
    const uint32_t base = (tma::smem_u32(smem_raw) + 1023u) & ~1023u;
    const uint32_t bar = base + 16384;
    if (threadIdx.x == 0) { tma::mbar_init(bar, 1); tma::fence_barrier_init(); }
    __syncthreads();
    if (threadIdx.x == 0) {
        tma::prefetch_tmap(&map);
        tma::mbar_expect_tx(bar, 2048 * boxes);
        for (int b = 0; b < boxes; ++b) tma::load_2d(base + 2048 * b, &map, bar, 0, (int)blockIdx.x * 16 * boxes + 16 * b);
    }
    tma::mbar_wait(bar, 0);
    double v;
    asm volatile("ld.shared.f64 %0, [%1];" : "=d"(v) : "r"(base + 8 * threadIdx.x));
    if (v == 12345.678) out[0] = v;
}
// first data through plain loads: same bytes, 128-bit LDG + st.shared + __syncthreads
__global__ void k_ldg_first(const double *src, double *out, int boxes, int ld)
{
    extern __shared__ uint8_t smem_raw[];
    double *s = reinterpret_cast<double *>(smem_raw);
    const int rows = 16 * boxes;
    for (int i = threadIdx.x; i < rows * 8; i += blockDim.x) {
        const int r = i >> 3, c = i & 7;
        const double2 v = *reinterpret_cast<const double2 *>(src + (size_t)(blockIdx.x * rows + r) * ld + 2 * c);
        *reinterpret_cast<double2 *>(s + r * 16 + 2 * c) = v;
    }
    __syncthreads();
    if (s[threadIdx.x] == 12345.678) out[0] = s[threadIdx.x];
}

template <typename F> float per_launch_us(F f, int iters = 400)
{
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    float best = 1e30f;
    for (int rep = 0; rep < 4; ++rep) {
        cudaDeviceSynchronize();
        cudaEventRecord(e0);
        for (int i = 0; i < iters; ++i) f();
        cudaEventRecord(e1);
        cudaDeviceSynchronize();
        float t; cudaEventElapsedTime(&t, e0, e1);
        best = std::min(best, t * 1e3f / iters);
    }
    return best;
}

int main()
{
    double *src, *out; int *flag;
    const int ld = 512, rows = 1 << 16;
    cudaMalloc(&src, sizeof(double) * ld * rows); cudaMemset(src, 0, sizeof(double) * ld * rows);
    cudaMalloc(&out, 64); cudaMalloc(&flag, 64);
    CUtensorMap map;
    if (!tma::make_2d(&map, CU_TENSOR_MAP_DATA_TYPE_FLOAT64, 8, src, rows, ld, 16, 16)) { printf("map failed\n"); return 1; }
    cudaFuncSetAttribute(k_dynsmem, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
    cudaFuncSetAttribute(k_tma_first, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
    cudaFuncSetAttribute(k_ldg_first, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
    Big big{};
    for (int grid : {1, 64, 148, 256, 1036}) {
        printf("{\"grid\":%d", grid);
        printf(",\"empty_160\":%.2f", per_launch_us([&] { k_empty<<<grid, 160>>>(flag); }));
        printf(",\"empty_288\":%.2f", per_launch_us([&] { k_empty<<<grid, 288>>>(flag); }));
        printf(",\"dynsmem_25k\":%.2f", per_launch_us([&] { k_dynsmem<<<grid, 160, 25 * 1024>>>(flag); }));
        printf(",\"dynsmem_100k\":%.2f", per_launch_us([&] { k_dynsmem<<<grid, 160, 100 * 1024>>>(flag); }));
        printf(",\"dynsmem_197k\":%.2f", per_launch_us([&] { k_dynsmem<<<grid, 288, 197 * 1024>>>(flag); }));
        printf(",\"bigparam\":%.2f", per_launch_us([&] { k_bigparam<<<grid, 160>>>(big, flag); }));
        for (int boxes : {1, 4})
            printf(",\"tma_first_%d\":%.2f", boxes, per_launch_us([&] { k_tma_first<<<grid, 160, 25 * 1024>>>(map, out, boxes); }));
        for (int boxes : {1, 4})
            printf(",\"ldg_first_%d\":%.2f", boxes, per_launch_us([&] { k_ldg_first<<<grid, 160, 25 * 1024>>>(src, out, boxes, ld); }));
        printf("}\n");
    }
    return 0;
}
