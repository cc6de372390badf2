// tma.cuh -- mbarrier / TMA (cp.async.bulk.tensor) PTX wrappers and host-side tensor-map
// encoding for the sm_100a GEMM kernels.  The driver entry point is resolved at run time
// through the runtime API, so the library carries no link-time dependency on libcuda.
#pragma once
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <cstdlib>

namespace accblas {
namespace tma {

__device__ __forceinline__ uint32_t smem_u32(const void *p) {
    return static_cast<uint32_t>(__cvta_generic_to_shared(p));
}
__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count));
}
__device__ __forceinline__ void fence_barrier_init() {
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void fence_proxy_async() {
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
}
// Before a consumer warp hands a stage back to the TMA producer: the warp's ld.shared reads of the
// stage (generic proxy) must be COMPLETE before the producer's next cp.async.bulk (async proxy) may
// overwrite it.  The mbarrier release/acquire chain alone does not order accesses across the two
// proxies, and ptxas is free to hoist the arrive above the MMAs that consume the loaded registers
// (it did, in the 32 x 32 complex tile: SYNCS.ARRIVE right behind the last LDS.128, wrong results
// under load).  The proxy fence orders the reads; __syncwarp extends it to the whole warp.
__device__ __forceinline__ void consumer_release_fence() {
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    __syncwarp();
}
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint32_t bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ bool mbar_try_wait(uint32_t bar, uint32_t parity) {
    uint32_t ok;
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(ok) : "r"(bar), "r"(parity) : "memory");
    return ok != 0;
}
// Spin on try_wait (each probe already suspends the thread for a hardware time slice).  A wait that
// has not completed after ~2^26 probes is a protocol bug: trap instead of hanging the GPU.
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
    uint32_t spins = 0;
    while (!mbar_try_wait(bar, parity)) {
        if (++spins > (1u << 26)) __trap();
    }
}
__device__ __forceinline__ void prefetch_tmap(const void *tmap) {
    asm volatile("prefetch.tensormap [%0];" ::"l"(tmap) : "memory");
}
// 2-D tiled load: box lands at `dst` (shared), completion bytes are posted on `bar`.
__device__ __forceinline__ void load_2d(uint32_t dst, const void *tmap, uint32_t bar, int c0, int c1) {
    asm volatile(
        "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
        ::"r"(dst), "l"(tmap), "r"(bar), "r"(c0), "r"(c1) : "memory");
}

// 3-D tiled load (an MN-major operand panel as ONE box: see make_mn3d)
__device__ __forceinline__ void load_3d(uint32_t dst, const void *tmap, uint32_t bar, int c0, int c1, int c2) {
    asm volatile(
        "cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5}], [%2];"
        ::"r"(dst), "l"(tmap), "r"(bar), "r"(c0), "r"(c1), "r"(c2) : "memory");
}
// box (c0, c1) of plane `plane` of a strided batch of matrices (make_2d_batched), or of a single matrix (plane < 0,
// make_2d): what the batched launches of the gemm families use for every operand box
__device__ __forceinline__ void load_box(uint32_t dst, const void *tmap, uint32_t bar, int c0, int c1, int plane) {
    if (plane < 0) load_2d(dst, tmap, bar, c0, c1);
    else load_3d(dst, tmap, bar, c0, c1, plane);
}

// L2 eviction policies for TMA loads.  In a grouped rasterisation the row panels of A are re-read by
// every wave of the group while a panel of B is used by one wave and is then dead: keeping A
// (evict_last) and letting B go first (evict_first) turns part of the per-wave A re-reads into L2 hits.
__device__ __forceinline__ uint64_t policy_evict_last() {
    uint64_t p;
    asm volatile("createpolicy.fractional.L2::evict_last.b64 %0, 1.0;" : "=l"(p));
    return p;
}
__device__ __forceinline__ uint64_t policy_evict_first() {
    uint64_t p;
    asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(p));
    return p;
}
__device__ __forceinline__ void load_2d_hint(uint32_t dst, const void *tmap, uint32_t bar, int c0, int c1, uint64_t policy) {
    asm volatile(
        "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint [%0], [%1, {%3, %4}], [%2], %5;"
        ::"r"(dst), "l"(tmap), "r"(bar), "r"(c0), "r"(c1), "l"(policy) : "memory");
}

// ---------------------------------------------------------------------------------
// host side
// ---------------------------------------------------------------------------------
typedef CUresult (*EncodeTiledFn)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *,
                                  const cuuint64_t *, const cuuint32_t *, const cuuint32_t *,
                                  CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion,
                                  CUtensorMapFloatOOBfill);

inline EncodeTiledFn encode_fn()
{
    static EncodeTiledFn fn = [] {
        void *p = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) != cudaSuccess ||
            q != cudaDriverEntryPointSuccess)
            p = nullptr;
        return reinterpret_cast<EncodeTiledFn>(p);
    }();
    return fn;
}

// Encoding a tensor map costs the host about a microsecond -- a visible share of a 6 us call on the reference's small
// benchmark shapes, where a loop multiplies the same arrays again and again.  The map is a pure function of
// (type, rank, base, dims, strides, box), so a small per-thread direct-mapped cache of encoded maps is safe: a hit
// returns the bytes the driver produced for exactly these arguments (no device state is involved).
// ACCBLAS_TMAP_CACHE=0 disables it (A/B measurements).
struct MapKey {
    uint64_t base, dims[3], strides[2];
    uint32_t box[3], dt, rank;
    bool operator==(const MapKey &o) const {
        return base == o.base && dims[0] == o.dims[0] && dims[1] == o.dims[1] && dims[2] == o.dims[2] &&
               strides[0] == o.strides[0] && strides[1] == o.strides[1] && box[0] == o.box[0] && box[1] == o.box[1] &&
               box[2] == o.box[2] && dt == o.dt && rank == o.rank;
    }
};
inline bool encode_cached(CUtensorMap *map, CUtensorMapDataType dt, uint32_t rank, const void *base, const cuuint64_t *dims,
                          const cuuint64_t *strides, const cuuint32_t *box)
{
    EncodeTiledFn fn = encode_fn();
    if (!fn) return false;
    static const bool enabled = [] { const char *e = getenv("ACCBLAS_TMAP_CACHE"); return !(e && *e == '0'); }();
    constexpr int SLOTS = 64;
    struct Entry { MapKey key; CUtensorMap map; bool valid; };
    static thread_local Entry cache[SLOTS];
    MapKey k{};
    k.base = reinterpret_cast<uint64_t>(base); k.dt = (uint32_t)dt; k.rank = rank;
    for (uint32_t i = 0; i < rank; ++i) { k.dims[i] = dims[i]; k.box[i] = box[i]; }
    for (uint32_t i = 0; i + 1 < rank; ++i) k.strides[i] = strides[i];
    uint64_t h = k.base * 0x9E3779B97F4A7C15ull ^ (k.dims[0] * 31 + k.dims[1] * 131 + k.box[0] * 7 + k.box[1] * 3 + k.box[2] + rank);
    Entry &e = cache[(h >> 40) % SLOTS];
    if (enabled && e.valid && e.key == k) { *map = e.map; return true; }
    cuuint32_t estr[3] = {1, 1, 1};
    if (fn(map, dt, rank, const_cast<void *>(base), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
           CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) != CUDA_SUCCESS)
        return false;
    e.key = k; e.map = *map; e.valid = true;
    return true;
}

// Row-major rows x cols matrix of `esize`-byte elements viewed as a 2-D tensor
// (dim0 = columns, contiguous).  Box = box_cols x box_rows, 128-byte swizzle, zero fill
// outside the matrix.  Requires: base 16-byte aligned, cols * esize % 16 == 0,
// box_cols * esize <= 128.
inline bool make_2d(CUtensorMap *map, CUtensorMapDataType dt, size_t esize, const void *base, uint64_t rows,
                    uint64_t cols, uint32_t box_rows, uint32_t box_cols, uint64_t pitch_bytes = 0)
{
    cuuint64_t dims[2] = {cols, rows};
    cuuint64_t strides[1] = {pitch_bytes ? pitch_bytes : cols * esize};
    cuuint32_t box[2] = {box_cols, box_rows};
    return encode_cached(map, dt, 2, base, dims, strides, box);
}

// `planes` row-major rows x cols matrices, plane b starting plane_stride_bytes * b after `base` (a strided batch), as
// a 3-D tensor (dim2 = plane).  Box = box_cols x box_rows x 1: lands exactly like the 2-D box of make_2d, zero fill
// outside EACH matrix.  plane_stride_bytes must be a non-zero multiple of 16 (a batch that re-uses one operand for every
// product passes planes = 1 and always loads plane 0).
inline bool make_2d_batched(CUtensorMap *map, CUtensorMapDataType dt, size_t esize, const void *base, uint64_t rows,
                            uint64_t cols, uint32_t box_rows, uint32_t box_cols, uint64_t pitch_bytes, uint64_t planes,
                            uint64_t plane_stride_bytes)
{
    if (planes == 0 || plane_stride_bytes == 0 || (plane_stride_bytes & 15u)) return false;
    cuuint64_t dims[3] = {cols, rows, planes};
    cuuint64_t strides[2] = {pitch_bytes ? pitch_bytes : cols * esize, plane_stride_bytes};
    cuuint32_t box[3] = {box_cols, box_rows, 1};
    return encode_cached(map, dt, 3, base, dims, strides, box);
}

// An MN-major operand (rows = K, cols = the m or n extent, `cols` a multiple of `inner`) viewed as a 3-D tensor
//   dim0 = `inner` consecutive columns (one 128-byte swizzle row), dim1 = the K rows, dim2 = column group,
// so that ONE box {inner, box_k, groups} lands as `groups` consecutive [box_k][inner] sub-tiles -- exactly the
// shared-memory layout the kernels read when the panel is fetched as `groups` separate 2-D boxes, for one TMA
// instruction instead of eight.  dim2's stride (128 bytes) is smaller than dim1's (the row pitch): if the driver
// refuses such a map the caller keeps the 2-D boxes.
inline bool make_mn3d(CUtensorMap *map, CUtensorMapDataType dt, size_t esize, const void *base, uint64_t rows_k,
                      uint64_t cols, uint32_t inner, uint32_t box_k, uint32_t groups, uint64_t pitch_bytes)
{
    if (cols % inner != 0) return false;
    cuuint64_t dims[3] = {inner, rows_k, cols / inner};
    cuuint64_t strides[2] = {pitch_bytes, (cuuint64_t)inner * esize};
    cuuint32_t box[3] = {inner, box_k, groups};
    return encode_cached(map, dt, 3, base, dims, strides, box);
}

}  // namespace tma
}  // namespace accblas
