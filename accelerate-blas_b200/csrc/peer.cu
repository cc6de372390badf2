// peer.cu -- NVLink peer-memory plumbing for the one-process-per-GPU sharded products.
//
// The reference has no multi-GPU path at all (one cuBLAS handle per CUDA
// context, LLVM/PTX/Context.hs:50-65, is its only provision).  BASELINE.json configs[4] asks for
// rows of A split, B broadcast and C gathered.  On an NVSwitch box the gather needs no collective
// and no extra kernel: every rank maps the root's C through CUDA IPC and its gemm epilogue stores
// the finished tiles straight into the root's HBM over NVLink (plain st.global on a peer-mapped
// address), so the transfer overlaps the math tile by tile.  The A row panels travel the other
// way on the copy engines (cudaMemcpyAsync on a peer-mapped source), which costs the root no SMs.
//
//   accblas_peer_export   root:  describe the allocation behind a device pointer (IPC handle)
//   accblas_peer_open     peers: map it, get the pointer that corresponds to the root's
//   accblas_peer_copy2d   peers: pitched device-to-device copy on the caller's stream
//   accblas_peer_close_all
//
// Mappings are cached per process (opening a handle costs milliseconds for a multi-GiB
// allocation); this cache and the launch counters are the library's only process state.
#include "common.cuh"
#include <cuda.h>
#include <cstring>
#include <map>
#include <mutex>
#include <string>

namespace accblas {
namespace {

struct Mapping { void *base; };
std::mutex g_peer_mu;
std::map<std::string, Mapping> g_peer_map;       // key = the 64 handle bytes

}  // namespace
}  // namespace accblas

using namespace accblas;

extern "C" {

int accblas_peer_export(const void *dev_ptr, accblas_peer_handle_t *out)
{
    if (!dev_ptr || !out) return ACCBLAS_ERR_NULL;
    static_assert(sizeof(cudaIpcMemHandle_t) == 64, "IPC handle is 64 bytes");
    CUdeviceptr base = 0;
    size_t size = 0;
    // the IPC handle names the whole cudaMalloc allocation: remember where dev_ptr sits in it
    // (driver entry point resolved at run time: the library must load on a box without libcuda)
    typedef CUresult (*RangeFn)(CUdeviceptr *, size_t *, CUdeviceptr);
    static RangeFn range_fn = [] {
        void *p = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuMemGetAddressRange", &p, cudaEnableDefault, &q) != cudaSuccess ||
            q != cudaDriverEntryPointSuccess)
            p = nullptr;
        return reinterpret_cast<RangeFn>(p);
    }();
    if (!range_fn || range_fn(&base, &size, (CUdeviceptr)(uintptr_t)dev_ptr) != CUDA_SUCCESS)
        return cuda_fail(cudaErrorInvalidValue, "cuMemGetAddressRange");
    cudaIpcMemHandle_t h;
    ACC_CUDA(cudaIpcGetMemHandle(&h, (void *)(uintptr_t)base));
    memcpy(out->ipc, &h, 64);
    out->offset = (uint64_t)((uintptr_t)dev_ptr - (uintptr_t)base);
    out->size = (uint64_t)size;
    int dev = 0;
    ACC_CUDA(cudaGetDevice(&dev));
    out->device = dev;
    out->reserved = 0;
    return ACCBLAS_OK;
}

int accblas_peer_open(const accblas_peer_handle_t *h, void **mapped)
{
    if (!h || !mapped) return ACCBLAS_ERR_NULL;
    const std::string key(reinterpret_cast<const char *>(h->ipc), 64);
    std::lock_guard<std::mutex> lock(g_peer_mu);
    auto it = g_peer_map.find(key);
    if (it == g_peer_map.end()) {
        cudaIpcMemHandle_t ih;
        memcpy(&ih, h->ipc, 64);
        void *base = nullptr;
        ACC_CUDA(cudaIpcOpenMemHandle(&base, ih, cudaIpcMemLazyEnablePeerAccess));
        it = g_peer_map.emplace(key, Mapping{base}).first;
    }
    *mapped = static_cast<char *>(it->second.base) + h->offset;
    return ACCBLAS_OK;
}

int accblas_peer_close_all(void)
{
    std::lock_guard<std::mutex> lock(g_peer_mu);
    cudaError_t first = cudaSuccess;
    for (auto &kv : g_peer_map) {
        cudaError_t e = cudaIpcCloseMemHandle(kv.second.base);
        if (first == cudaSuccess) first = e;
    }
    g_peer_map.clear();
    return first == cudaSuccess ? ACCBLAS_OK : cuda_fail(first, "cudaIpcCloseMemHandle");
}

int accblas_peer_copy2d(accblas_stream_t stream, void *dst, size_t dst_pitch, const void *src, size_t src_pitch,
                        size_t width_bytes, size_t rows)
{
    if (width_bytes == 0 || rows == 0) return ACCBLAS_OK;
    if (!dst || !src) return ACCBLAS_ERR_NULL;
    if (dst_pitch < width_bytes || src_pitch < width_bytes) return ACCBLAS_ERR_SHAPE;
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    if (dst_pitch == width_bytes && src_pitch == width_bytes)
        ACC_CUDA(cudaMemcpyAsync(dst, src, width_bytes * rows, cudaMemcpyDefault, s));
    else
        ACC_CUDA(cudaMemcpy2DAsync(dst, dst_pitch, src, src_pitch, width_bytes, rows, cudaMemcpyDefault, s));
    return ACCBLAS_OK;
}

}  // extern "C"
