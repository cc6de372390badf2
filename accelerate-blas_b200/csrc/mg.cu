// mg.cu -- the multi-GPU entry points of the C ABI (accblas_mg_*): one process per GPU on an NVLink / NVSwitch box.
//
// NEW relative to the reference, which has no multi-GPU path (its only provision is one cuBLAS
// handle per CUDA context, LLVM/PTX/Context.hs:50-65).  BASELINE.json configs[4] / SURVEY.md 8(e):
//
//   gemm   rows of op(A) and C split into contiguous panels, one per rank; operands start and the result ends on
//          `root`.  Every other rank maps the root's A, B and C through CUDA IPC and
//            * PULLS its operands with the copy engines, K-chunk by K-chunk (a column slab of its A panel and the
//              matching rows of B per chunk) -- no SMs on either side, no collective kernel to schedule;
//            * multiplies chunk c as soon as it has landed (stream events), accumulating into a LOCAL panel through
//              the beta = 1 epilogue (accblas_gemm_ex), so the transfers of chunks c+1.. overlap the product of chunk c;
//              only the first, small chunk is exposed;
//            * runs the LAST chunk with C pointing INTO THE ROOT'S HBM: the epilogue's stores are the gather.
//          NCCL carries two 8-byte tokens per call: "operands ready" (orders the peers' pulls after the root's
//          preceding work on its stream) and "done" (orders the root's stream after every peer's remote stores).
//   gemv   op N: rows of A and y split (A pre-sharded), x replicated, y slices ALL-GATHERED;
//          op T/H: rows of A and x split, every rank forms a full-length partial y, ALL-REDUCED in rank order.
//   dot    contiguous slices, the partial sums ALL-REDUCED in rank order.
//          The gemv / dot exchanges are a few KiB: they go through a small IPC-mapped exchange buffer per rank --
//          every rank stores its piece straight into every peer's buffer, raises a flag (st.release.sys) and waits
//          for the peers' flags -- one tiny kernel instead of a 40-60 us NCCL collective.  Sums run in rank order:
//          bit-reproducible, identical on every rank.
//
// NCCL is loaded with dlopen at accblas_mg_init (libnccl.so.2, or $ACCBLAS_NCCL_LIB): the library keeps its
// cudart-only link line, and single-GPU users never touch NCCL.  The bootstrap id is plain bytes the caller ships
// with any transport (torch.distributed, MPI, a pipe), like the peer handle of accblas_peer_*.
#include "common.cuh"
#include "mg_exchange.cuh"
#include <dlfcn.h>
#include <algorithm>
#include <cstdio>
#include <cstring>
#include <vector>

namespace accblas {
namespace {

// ------------------------------------------------------------------------------------------------------------
// NCCL through dlopen: the handful of prototypes used, restated (nccl.h 2.27 / 2.28 agree on all of them)
// ------------------------------------------------------------------------------------------------------------
struct NcclId { char internal[128]; };           // ncclUniqueId, passed BY VALUE to ncclCommInitRank
typedef void *NcclComm;
enum { NCCL_UINT8 = 1, NCCL_FLOAT64 = 8, NCCL_SUM = 0 };
struct Nccl {
    void *h = nullptr;
    int (*GetUniqueId)(NcclId *) = nullptr;
    int (*CommInitRank)(NcclComm *, int, NcclId, int) = nullptr;
    int (*CommDestroy)(NcclComm) = nullptr;
    int (*Broadcast)(const void *, void *, size_t, int, int, NcclComm, cudaStream_t) = nullptr;
    int (*AllReduce)(const void *, void *, size_t, int, int, NcclComm, cudaStream_t) = nullptr;
    int (*AllGather)(const void *, void *, size_t, int, NcclComm, cudaStream_t) = nullptr;
    const char *(*GetErrorString)(int) = nullptr;
    char why[256] = {0};
    bool load()
    {
        if (h) return true;
        const char *names[3] = {getenv("ACCBLAS_NCCL_LIB"), "libnccl.so.2", "libnccl.so"};
        for (const char *n : names) {
            if (!n || !*n) continue;
            h = dlopen(n, RTLD_NOW | RTLD_GLOBAL);
            if (h) break;
            snprintf(why, sizeof why, "%s", dlerror());
        }
        if (!h) return false;
#define SYM(f) *(void **)(&f) = dlsym(h, "nccl" #f); if (!f) { snprintf(why, sizeof why, "nccl" #f " not found"); return false; }
        SYM(GetUniqueId) SYM(CommInitRank) SYM(CommDestroy) SYM(Broadcast) SYM(AllReduce) SYM(AllGather) SYM(GetErrorString)
#undef SYM
        return true;
    }
};
Nccl g_nccl;

int nccl_fail(int rc, const char *where)
{
    char buf[200];
    snprintf(buf, sizeof buf, "%s: NCCL error %d (%s)", where, rc, g_nccl.GetErrorString ? g_nccl.GetErrorString(rc) : "?");
    return cuda_fail(cudaErrorUnknown, buf);
}
#define ACC_NCCL(call) do { int rc_ = (call); if (rc_ != 0) return nccl_fail(rc_, #call); } while (0)

// ------------------------------------------------------------------------------------------------------------
// communicator
// ------------------------------------------------------------------------------------------------------------
struct Comm {
    int rank = 0, world = 1, dev = 0;
    NcclComm nccl = nullptr;
    cudaStream_t pull = nullptr;                   // copy-engine pulls of the gemm operands
    double *token = nullptr;                       // device scalar for the ready / done tokens
    // exchange buffer (IPC-mapped into every peer): [parity 0 | parity 1], each = flags + world * XSLOT
    char *xbuf = nullptr;
    char *peer_x[MG_MAX_RANKS] = {nullptr};        // peer_x[r]: rank r's buffer as mapped here (own buffer for r == rank)
    size_t xparity = 0;                            // bytes per parity
    uint32_t seq = 0;                              // exchange collectives issued so far
    // sharded dot (one launch, dot.cu): [arrival counter | per-CTA partials]; the counter returns to zero by itself
    char *dot_scratch = nullptr;
    // staging for the plan's handle broadcast: owned by the communicator so that a plan never has to allocate
    // (and possibly fail) between deciding to enter a collective and entering it
    char *wire = nullptr;
};
constexpr size_t MG_WIRE_BYTES = 512;
constexpr int MG_DOT_GRID = 1024;                  // CTAs the fused dot may launch (partials: 32 bytes each)

// ------------------------------------------------------------------------------------------------------------
// exchange kernels (protocol and flag helpers: mg_exchange.cuh)
// ------------------------------------------------------------------------------------------------------------
// All-reduce of ONE element (dot): thread r deposits this rank's value in rank r's buffer and raises the flag;
// then waits for rank r's deposit here; thread 0 adds the deposits in rank order.
template <typename T>
__global__ void __launch_bounds__(32) mg_scalar_allreduce_kernel(PeerPtrs peers, int rank, int world, size_t parity_off,
                                                                  uint32_t seq, const T *mine, T *out)
{
    const int r = threadIdx.x;
    __shared__ T got[MG_MAX_RANKS];
    if (r < world) {
        char *dst = peers.p[r] + parity_off;
        *reinterpret_cast<T *>(dst + XFLAGS + (size_t)rank * XSLOT) = *mine;
        __threadfence_system();
        st_release_sys(reinterpret_cast<uint32_t *>(dst + 64 * rank), seq);
        const char *own = peers.p[rank] + parity_off;
        wait_flag(reinterpret_cast<const uint32_t *>(own + 64 * r), seq);
        got[r] = __ldcv(reinterpret_cast<const T *>(own + XFLAGS + (size_t)r * XSLOT));
    }
    __syncthreads();
    if (r == 0) {
        T acc = got[0];
        for (int i = 1; i < world; ++i) acc = add_s(acc, got[i]);
        *out = acc;
    }
}

// Deposit `bytes` (a multiple of 16, 16-byte aligned) of `src` in every peer's slot for this rank.  grid = (blocks, world).
__global__ void __launch_bounds__(256) mg_push_kernel(PeerPtrs peers, int rank, size_t parity_off, const int4 *src, size_t vecs)
{
    int4 *dst = reinterpret_cast<int4 *>(peers.p[blockIdx.y] + parity_off + XFLAGS + (size_t)rank * XSLOT);
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < vecs; i += (size_t)gridDim.x * blockDim.x) dst[i] = src[i];
}

// After mg_push_kernel has COMPLETED (kernel boundary: its remote stores are performed): CTA 0 raises this rank's
// flag in every peer; every CTA waits for all peers' flags here, then
//   GATHER : copies slice r (slice_bytes each) of the result from rank r's deposit            (gemv N)
//   REDUCE : sums the deposits elementwise in rank order                                       (gemv T/H)
// The deposits of a GATHER need not have the same length (accblas_mg_row_panel hands out whole 128-row tiles: m = 1000
// on 8 ranks is 7 x 128 + 104, m = 256 leaves ranks 2..7 empty): every rank writes its element count into the spare
// bytes of its 64-byte flag line (offset 8) before it raises the flag, and the gather places deposit r behind the
// deposits of ranks 0..r-1.
template <typename T, bool REDUCE>
__global__ void __launch_bounds__(256) mg_collect_kernel(PeerPtrs peers, int rank, int world, size_t parity_off, uint32_t seq,
                                                         T *out, size_t elems /*this rank's deposit*/)
{
    if (blockIdx.x == 0 && threadIdx.x < world) {
        char *line = peers.p[threadIdx.x] + parity_off + 64 * rank;
        *reinterpret_cast<unsigned long long *>(line + 8) = (unsigned long long)elems;
        __threadfence_system();
        st_release_sys(reinterpret_cast<uint32_t *>(line), seq);
    }
    const char *own = peers.p[rank] + parity_off;
    __shared__ unsigned long long count[MG_MAX_RANKS];
    if (threadIdx.x < world) {
        wait_flag(reinterpret_cast<const uint32_t *>(own + 64 * threadIdx.x), seq);
        count[threadIdx.x] = __ldcv(reinterpret_cast<const unsigned long long *>(own + 64 * threadIdx.x + 8));
    }
    __syncthreads();
    const size_t stride = (size_t)gridDim.x * blockDim.x, i0 = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (REDUCE) {
        for (size_t i = i0; i < elems; i += stride) {
            T acc = __ldcv(reinterpret_cast<const T *>(own + XFLAGS) + i);
            for (int r = 1; r < world; ++r) acc = add_s(acc, __ldcv(reinterpret_cast<const T *>(own + XFLAGS + (size_t)r * XSLOT) + i));
            out[i] = acc;
        }
    } else {
        size_t at = 0;
        for (int r = 0; r < world; ++r) {
            const size_t len = (size_t)count[r];
            for (size_t i = i0; i < len; i += stride)
                out[at + i] = __ldcv(reinterpret_cast<const T *>(own + XFLAGS + (size_t)r * XSLOT) + i);
            at += len;
        }
    }
}

PeerPtrs peer_ptrs(const Comm *c)
{
    PeerPtrs pp;
    for (int r = 0; r < MG_MAX_RANKS; ++r) pp.p[r] = c->peer_x[r < c->world ? r : c->rank];
    return pp;
}

// deposit + collect of `bytes` per rank on stream s (shared by gemv N / T / H)
template <typename T, bool REDUCE>
int exchange(Comm *c, cudaStream_t s, const void *mine, size_t elems, T *out)
{
    const size_t bytes = elems * sizeof(T), vecs = (bytes + 15) / 16;
    const uint32_t seq = ++c->seq;
    const size_t off = (seq & 1u) * c->xparity;
    const PeerPtrs pp = peer_ptrs(c);
    dim3 grid((unsigned)std::max<size_t>(1, std::min<size_t>((vecs + 255) / 256, 16)), (unsigned)c->world);
    mg_push_kernel<<<grid, 256, 0, s>>>(pp, c->rank, off, static_cast<const int4 *>(mine), vecs);
    ACC_TRY(check_launch("mg_push_kernel"));
    const size_t total = REDUCE ? elems : std::max<size_t>(elems, 256) * (size_t)c->world;   // gather: the peers' lengths are not known here
    const unsigned blocks = (unsigned)std::max<size_t>(1, std::min<size_t>((total + 255) / 256, 2 * (size_t)sm_count()));
    mg_collect_kernel<T, REDUCE><<<blocks, 256, 0, s>>>(pp, c->rank, c->world, off, seq, out, elems);
    return check_launch("mg_collect_kernel");
}

// ------------------------------------------------------------------------------------------------------------
// gemm plan
// ------------------------------------------------------------------------------------------------------------
struct Chunk { int64_t k0, dk; };

struct Plan {
    Comm *comm = nullptr;
    int dtype = 0, opA = 0, opB = 0, root = 0;
    int64_t rowsA = 0, colsA = 0, rowsB = 0, colsB = 0, m = 0, n = 0, k = 0;
    int64_t start = 0, rows = 0;                  // this rank's panel of op(A) / C
    std::vector<Chunk> chunks;
    // root: its own operands; peers: the root's arrays as mapped here
    const char *A = nullptr, *B = nullptr;
    char *C = nullptr;
    // peers: local replicas
    char *A_loc = nullptr, *B_loc = nullptr, *C_acc = nullptr;
    std::vector<cudaEvent_t> landed;              // chunk c's operands are in local memory
    cudaEvent_t go = nullptr;
    bool as_peer = false;                         // this rank runs the pull / chunk pipeline (every rank but the root;
                                                  // the root too under ACCBLAS_TUNE_MG_SELF_PEER, a one-GPU test hook)
};

}  // namespace
}  // namespace accblas

using namespace accblas;

extern "C" {

int accblas_mg_unique_id(void *id_out)
{
    if (!id_out) return ACCBLAS_ERR_NULL;
    if (!g_nccl.load()) return cuda_fail(cudaErrorNotSupported, g_nccl.why);
    NcclId id;
    ACC_NCCL(g_nccl.GetUniqueId(&id));
    memcpy(id_out, &id, sizeof id);
    return ACCBLAS_OK;
}

int accblas_mg_init(accblas_mg_comm_t *comm_out, int rank, int world, const void *id_bytes)
{
    if (!comm_out || !id_bytes) return ACCBLAS_ERR_NULL;
    if (world < 1 || world > MG_MAX_RANKS || rank < 0 || rank >= world) return ACCBLAS_ERR_SHAPE;
    ACC_TRY(device_check());
    if (!g_nccl.load()) return cuda_fail(cudaErrorNotSupported, g_nccl.why);
    Comm *c = new Comm;
    c->rank = rank; c->world = world;
    ACC_CUDA(cudaGetDevice(&c->dev));
    NcclId id;
    memcpy(&id, id_bytes, sizeof id);
    ACC_NCCL(g_nccl.CommInitRank(&c->nccl, world, id, rank));
    int lo = 0, hi = 0;
    cudaDeviceGetStreamPriorityRange(&lo, &hi);
    ACC_CUDA(cudaStreamCreateWithPriority(&c->pull, cudaStreamNonBlocking, hi));
    ACC_CUDA(cudaMalloc(reinterpret_cast<void **>(&c->token), 64));
    ACC_CUDA(cudaMemset(c->token, 0, 64));
    ACC_CUDA(cudaMalloc(reinterpret_cast<void **>(&c->dot_scratch), 64 + 32 * (size_t)MG_DOT_GRID));
    ACC_CUDA(cudaMemset(c->dot_scratch, 0, 64 + 32 * (size_t)MG_DOT_GRID));
    ACC_CUDA(cudaMalloc(reinterpret_cast<void **>(&c->wire), MG_WIRE_BYTES));
    // exchange buffer: cudaMalloc (IPC needs a plain allocation), zeroed, handle all-gathered, peers mapped
    c->xparity = XFLAGS + (size_t)world * XSLOT;
    ACC_CUDA(cudaMalloc(reinterpret_cast<void **>(&c->xbuf), 2 * c->xparity));
    ACC_CUDA(cudaMemset(c->xbuf, 0, 2 * c->xparity));
    cudaIpcMemHandle_t mine;
    ACC_CUDA(cudaIpcGetMemHandle(&mine, c->xbuf));
    unsigned char *hd = nullptr;
    ACC_CUDA(cudaMalloc(reinterpret_cast<void **>(&hd), 64 * (size_t)(world + 1)));
    ACC_CUDA(cudaMemcpy(hd + 64 * (size_t)world, &mine, 64, cudaMemcpyHostToDevice));
    ACC_NCCL(g_nccl.AllGather(hd + 64 * (size_t)world, hd, 64, NCCL_UINT8, c->nccl, nullptr));
    ACC_CUDA(cudaStreamSynchronize(nullptr));
    std::vector<unsigned char> all(64 * (size_t)world);
    ACC_CUDA(cudaMemcpy(all.data(), hd, all.size(), cudaMemcpyDeviceToHost));
    cudaFree(hd);
    for (int r = 0; r < world; ++r) {
        if (r == rank) { c->peer_x[r] = c->xbuf; continue; }
        cudaIpcMemHandle_t h;
        memcpy(&h, all.data() + 64 * (size_t)r, 64);
        void *p = nullptr;
        ACC_CUDA(cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess));
        c->peer_x[r] = static_cast<char *>(p);
    }
    // nobody may deposit before everyone has zeroed and mapped: one all-reduce as the barrier
    ACC_NCCL(g_nccl.AllReduce(c->token, c->token, 1, NCCL_FLOAT64, NCCL_SUM, c->nccl, nullptr));
    ACC_CUDA(cudaStreamSynchronize(nullptr));
    *comm_out = reinterpret_cast<accblas_mg_comm_t>(c);
    return ACCBLAS_OK;
}

int accblas_mg_finalize(accblas_mg_comm_t comm)
{
    Comm *c = reinterpret_cast<Comm *>(comm);
    if (!c) return ACCBLAS_OK;
    cudaDeviceSynchronize();
    for (int r = 0; r < c->world; ++r)
        if (r != c->rank && c->peer_x[r]) cudaIpcCloseMemHandle(c->peer_x[r]);
    if (c->nccl) g_nccl.CommDestroy(c->nccl);
    if (c->xbuf) cudaFree(c->xbuf);
    if (c->token) cudaFree(c->token);
    if (c->dot_scratch) cudaFree(c->dot_scratch);
    if (c->wire) cudaFree(c->wire);
    if (c->pull) cudaStreamDestroy(c->pull);
    delete c;
    return ACCBLAS_OK;
}

int accblas_mg_rank(accblas_mg_comm_t comm) { return comm ? reinterpret_cast<Comm *>(comm)->rank : -1; }
int accblas_mg_world(accblas_mg_comm_t comm) { return comm ? reinterpret_cast<Comm *>(comm)->world : 0; }

// ---- partitioning, exported so that the host side (and the CPU tests) see exactly what the library does ----
int accblas_mg_row_panel(int64_t m, int world, int rank, int64_t *start, int64_t *rows)
{
    if (m < 0 || world < 1 || rank < 0 || rank >= world || !start || !rows) return ACCBLAS_ERR_SHAPE;
    int64_t per = (m + world - 1) / world;
    per = (per + 127) / 128 * 128;                 // whole CTA tiles per panel (keeps every window TMA-aligned)
    const int64_t s = std::min<int64_t>(m, per * rank);
    *start = s;
    *rows = std::min<int64_t>(per, m - s);
    return ACCBLAS_OK;
}

// K-chunks of the pipelined product: a small first chunk (the only transfer that is not hidden behind a product),
// then growing ones; all multiples of 512 so that every window keeps the TMA alignment.  Returns the count.
int accblas_mg_k_chunks(int64_t k, int64_t *k0_out, int64_t *dk_out, int cap)
{
    if (k < 0 || cap < 1 || !k0_out || !dk_out) return 0;
    static const int64_t ladder[4] = {512, 1536, 2048, 4096};
    int cnt = 0;
    int64_t pos = 0;
    if (k < 4096) { k0_out[0] = 0; dk_out[0] = k; return k > 0 ? 1 : 0; }
    while (pos < k && cnt < cap) {
        int64_t dk = ladder[std::min(cnt, 3)];
        if (cnt == cap - 1 || k - pos - dk < 512) dk = k - pos;      // the last chunk takes the rest
        k0_out[cnt] = pos; dk_out[cnt] = dk;
        pos += dk; ++cnt;
    }
    return cnt;
}

// ------------------------------------------------------------------------------------------------------------
// gemm: plan (collective; exchanges the IPC handles of the root's arrays, allocates the local replicas) / run / destroy
// ------------------------------------------------------------------------------------------------------------
int accblas_mg_gemm_plan(accblas_mg_comm_t comm, int dtype, int opA, int opB, int64_t rowsA, int64_t colsA, int64_t rowsB,
                         int64_t colsB, const void *A, const void *B, void *C, int root, accblas_mg_plan_t *plan_out)
{
    Comm *c = reinterpret_cast<Comm *>(comm);
    if (!c || !plan_out) return ACCBLAS_ERR_NULL;
    if (dtype < ACCBLAS_F32 || dtype > ACCBLAS_C64) return ACCBLAS_ERR_DTYPE;
    if (opA < 0 || opA > 2 || opB < 0 || opB > 2) return ACCBLAS_ERR_OP;
    if (rowsA < 0 || colsA < 0 || rowsB < 0 || colsB < 0 || root < 0 || root >= c->world) return ACCBLAS_ERR_SHAPE;
    Plan *p = new Plan;
    p->comm = c; p->dtype = dtype; p->opA = opA; p->opB = opB; p->root = root;
    p->rowsA = rowsA; p->colsA = colsA; p->rowsB = rowsB; p->colsB = colsB;
    p->m = opA == ACCBLAS_OP_N ? rowsA : colsA;
    p->k = opA == ACCBLAS_OP_N ? colsA : rowsA;
    p->n = opB == ACCBLAS_OP_N ? colsB : rowsB;
    if (p->k != (opB == ACCBLAS_OP_N ? rowsB : colsB)) { delete p; return ACCBLAS_ERR_SHAPE; }
    accblas_mg_row_panel(p->m, c->world, c->rank, &p->start, &p->rows);
    int64_t k0[64], dk[64];
    const int nc = accblas_mg_k_chunks(p->k, k0, dk, 64);
    for (int i = 0; i < nc; ++i) p->chunks.push_back(Chunk{k0[i], dk[i]});
    const size_t es = elem_size(dtype);
    // The root describes the allocations behind A, B, C; everyone else maps them (accblas_peer_*).  From here on the
    // call is COLLECTIVE: no rank may leave between the broadcast and the closing agreement, whatever failed locally,
    // or the others would wait in a collective forever.  Failures are recorded in `st` and travel with the data.
    struct Wire { accblas_peer_handle_t hs[3]; int32_t status; int32_t pad[3]; };
    static_assert(sizeof(Wire) <= MG_WIRE_BYTES, "the plan's broadcast fits the communicator's staging buffer");
    Wire w;
    memset(&w, 0, sizeof w);
    int st = ACCBLAS_OK;
    if (c->rank == root) {
        if ((p->m > 0 && p->n > 0 && !C) || (p->k > 0 && p->m > 0 && p->n > 0 && (!A || !B))) st = ACCBLAS_ERR_NULL;
        const void *ptrs[3] = {A, B, C};
        for (int i = 0; i < 3 && st == ACCBLAS_OK; ++i)
            if (ptrs[i] && c->world > 1) st = accblas_peer_export(ptrs[i], &w.hs[i]);
        p->A = static_cast<const char *>(A); p->B = static_cast<const char *>(B); p->C = static_cast<char *>(C);
        w.status = st;
    }
    if (c->world > 1) {
        // the root's verdict rides with the handles: a root that could not export gives EVERY rank the same error
        if (c->rank == root) {
            const cudaError_t e = cudaMemcpy(c->wire, &w, sizeof w, cudaMemcpyHostToDevice);
            if (e != cudaSuccess && st == ACCBLAS_OK) st = cuda_fail(e, "accblas_mg_gemm_plan: staging the handles");
        }
        const int rc = g_nccl.Broadcast(c->wire, c->wire, sizeof w, NCCL_UINT8, root, c->nccl, nullptr);
        if (rc != 0) { delete p; return nccl_fail(rc, "accblas_mg_gemm_plan: ncclBroadcast"); }   // the communicator itself is gone
        cudaError_t e = cudaStreamSynchronize(nullptr);
        if (e == cudaSuccess) e = cudaMemcpy(&w, c->wire, sizeof w, cudaMemcpyDeviceToHost);
        if (e != cudaSuccess && st == ACCBLAS_OK) st = cuda_fail(e, "accblas_mg_gemm_plan: reading the handles");
        if (st == ACCBLAS_OK && w.status != ACCBLAS_OK) {
            st = w.status;
            fail_text(st, "accblas_mg_gemm_plan: the root rank could not describe its operands (see its own error text)");
        }
    }
    const bool self_peer = c->rank == root && g_tuning[ACCBLAS_TUNE_MG_SELF_PEER] == 1;
    p->as_peer = c->rank != root || self_peer;
    auto local_setup = [&]() -> int {
        if (p->as_peer && p->rows > 0 && p->n > 0) {
            if (!self_peer) {
                void *mapped[3] = {nullptr, nullptr, nullptr};
                for (int i = 0; i < 3; ++i)
                    if (w.hs[i].size) ACC_TRY(accblas_peer_open(&w.hs[i], &mapped[i]));
                p->A = static_cast<const char *>(mapped[0]); p->B = static_cast<const char *>(mapped[1]); p->C = static_cast<char *>(mapped[2]);
            }
            // local replicas: this rank's panel of A (stored layout), all of B, and -- when the product is chunked --
            // the accumulator panel (the last chunk writes to the root instead)
            const size_t bytesA = es * (size_t)p->rows * (size_t)p->k, bytesB = es * (size_t)p->k * (size_t)p->n;
            if (p->k > 0) {
                ACC_CUDA(cudaMalloc(reinterpret_cast<void **>(&p->A_loc), std::max<size_t>(bytesA, 16)));
                ACC_CUDA(cudaMalloc(reinterpret_cast<void **>(&p->B_loc), std::max<size_t>(bytesB, 16)));
            }
            if (p->chunks.size() > 1) ACC_CUDA(cudaMalloc(reinterpret_cast<void **>(&p->C_acc), es * (size_t)p->rows * (size_t)p->n));
            for (size_t i = 0; i < p->chunks.size(); ++i) {
                cudaEvent_t e;
                ACC_CUDA(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
                p->landed.push_back(e);
            }
        }
        ACC_CUDA(cudaEventCreateWithFlags(&p->go, cudaEventDisableTiming));
        return ACCBLAS_OK;
    };
    if (st == ACCBLAS_OK) st = local_setup();
    if (c->world > 1) {
        // closing agreement: a plan exists on every rank or on none (a rank that could not map or allocate must not
        // leave the others with a plan whose run would wait for it).  The token carries the count of failed ranks.
        const double mine = st == ACCBLAS_OK ? 0.0 : 1.0;
        double failed = 1.0;
        cudaError_t e = cudaMemcpy(c->token, &mine, sizeof mine, cudaMemcpyHostToDevice);
        const int rc = g_nccl.AllReduce(c->token, c->token, 1, NCCL_FLOAT64, NCCL_SUM, c->nccl, nullptr);
        if (rc != 0) { accblas_mg_gemm_plan_destroy(reinterpret_cast<accblas_mg_plan_t>(p)); return nccl_fail(rc, "accblas_mg_gemm_plan: ncclAllReduce"); }
        if (e == cudaSuccess) e = cudaStreamSynchronize(nullptr);
        if (e == cudaSuccess) e = cudaMemcpy(&failed, c->token, sizeof failed, cudaMemcpyDeviceToHost);
        cudaMemset(c->token, 0, sizeof(double));                                  // the run tokens add zeros
        if (e != cudaSuccess && st == ACCBLAS_OK) st = cuda_fail(e, "accblas_mg_gemm_plan: closing agreement");
        if (st == ACCBLAS_OK && failed != 0.0)
            st = cuda_fail(cudaErrorUnknown, "accblas_mg_gemm_plan: another rank could not set its plan up (see that rank's error text)");
    }
    if (st != ACCBLAS_OK) {
        accblas_mg_gemm_plan_destroy(reinterpret_cast<accblas_mg_plan_t>(p));     // frees whatever was set up
        return st;
    }
    *plan_out = reinterpret_cast<accblas_mg_plan_t>(p);
    return ACCBLAS_OK;
}

int accblas_mg_gemm_plan_destroy(accblas_mg_plan_t plan)
{
    Plan *p = reinterpret_cast<Plan *>(plan);
    if (!p) return ACCBLAS_OK;
    cudaDeviceSynchronize();
    if (p->A_loc) cudaFree(p->A_loc);
    if (p->B_loc) cudaFree(p->B_loc);
    if (p->C_acc) cudaFree(p->C_acc);
    for (cudaEvent_t e : p->landed) cudaEventDestroy(e);
    if (p->go) cudaEventDestroy(p->go);
    delete p;
    return ACCBLAS_OK;
}

int accblas_mg_gemm_run(accblas_mg_plan_t plan, accblas_stream_t stream, const void *alpha_host)
{
    Plan *p = reinterpret_cast<Plan *>(plan);
    if (!p || !alpha_host) return ACCBLAS_ERR_NULL;
    Comm *c = p->comm;
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    const size_t es = elem_size(p->dtype);
    const bool an = p->opA == ACCBLAS_OP_N, bn = p->opB == ACCBLAS_OP_N;
    if (p->m == 0 || p->n == 0) return ACCBLAS_OK;
    // "operands ready": ordered after everything the root has enqueued on its stream so far
    if (c->world > 1) ACC_NCCL(g_nccl.AllReduce(c->token, c->token, 1, NCCL_FLOAT64, NCCL_SUM, c->nccl, s));

    static const double one_d[2] = {1.0, 0.0};
    static const float one_f[2] = {1.f, 0.f};
    const void *one = (p->dtype == ACCBLAS_F32 || p->dtype == ACCBLAS_C32) ? static_cast<const void *>(one_f) : one_d;

    // Everything between the two tokens is LOCAL work.  Its first failure ends the local work only: the "done" token
    // below is still enqueued, because the other ranks enter it whatever happened here (a rank that returned early
    // would leave them waiting in the all-reduce).
    auto local_work = [&]() -> int {
    int st = ACCBLAS_OK;
    if (!p->as_peer) {
        // the root multiplies its own panel in place: nothing to move
        if (p->rows > 0) {
            GemmArgs g{};
            g.dtype = p->dtype; g.opA = p->opA; g.opB = p->opB;
            g.rowsA = an ? p->rows : p->rowsA; g.colsA = an ? p->colsA : p->rows;
            g.rowsB = p->rowsB; g.colsB = p->colsB;
            g.lda = p->colsA; g.ldb = p->colsB; g.ldc = p->n;
            g.alpha = alpha_host;
            g.A = p->A + es * (size_t)(an ? p->start * p->colsA : p->start);
            g.B = p->B;
            g.C = p->C + es * (size_t)p->start * (size_t)p->n;
            g.stream = s;
            st = gemm_device(g);
        }
    } else if (p->rows > 0) {
        if (p->k == 0) {
            ACC_CUDA(cudaMemset2DAsync(p->C + es * (size_t)p->start * (size_t)p->n, es * (size_t)p->n, 0, es * (size_t)p->n, (size_t)p->rows, s));
        } else {
            ACC_CUDA(cudaEventRecord(p->go, s));
            ACC_CUDA(cudaStreamWaitEvent(c->pull, p->go, 0));
            const size_t nchunks = p->chunks.size();
            // ---- pulls: chunk by chunk, A slab then B slab, all on the copy engines ----
            for (size_t i = 0; i < nchunks; ++i) {
                const int64_t k0 = p->chunks[i].k0, dk = p->chunks[i].dk;
                if (an)   // stored m x k: rows [start, start+rows), columns [k0, k0+dk)
                    ACC_TRY(accblas_peer_copy2d(c->pull, p->A_loc + es * (size_t)k0, es * (size_t)p->k,
                                                p->A + es * ((size_t)p->start * (size_t)p->k + (size_t)k0), es * (size_t)p->k,
                                                es * (size_t)dk, (size_t)p->rows));
                else      // stored k x m: rows [k0, k0+dk), columns [start, start+rows)
                    ACC_TRY(accblas_peer_copy2d(c->pull, p->A_loc + es * (size_t)k0 * (size_t)p->rows, es * (size_t)p->rows,
                                                p->A + es * ((size_t)k0 * (size_t)p->m + (size_t)p->start), es * (size_t)p->m,
                                                es * (size_t)p->rows, (size_t)dk));
                if (bn)   // stored k x n: rows [k0, k0+dk) are contiguous
                    ACC_TRY(accblas_peer_copy2d(c->pull, p->B_loc + es * (size_t)k0 * (size_t)p->n, es * (size_t)p->n * (size_t)dk,
                                                p->B + es * (size_t)k0 * (size_t)p->n, es * (size_t)p->n * (size_t)dk,
                                                es * (size_t)p->n * (size_t)dk, 1));
                else      // stored n x k: columns [k0, k0+dk) of every row
                    ACC_TRY(accblas_peer_copy2d(c->pull, p->B_loc + es * (size_t)k0, es * (size_t)p->k,
                                                p->B + es * (size_t)k0, es * (size_t)p->k, es * (size_t)dk, (size_t)p->n));
                ACC_CUDA(cudaEventRecord(p->landed[i], c->pull));
            }
            // ---- products: chunk c accumulates into the local panel; the last one stores into the root's C ----
            char *c_root = p->C + es * (size_t)p->start * (size_t)p->n;
            for (size_t i = 0; i < nchunks && st == ACCBLAS_OK; ++i) {
                const int64_t k0 = p->chunks[i].k0, dk = p->chunks[i].dk;
                ACC_CUDA(cudaStreamWaitEvent(s, p->landed[i], 0));
                GemmArgs g{};
                g.dtype = p->dtype; g.opA = p->opA; g.opB = p->opB;
                g.rowsA = an ? p->rows : dk; g.colsA = an ? dk : p->rows;
                g.rowsB = bn ? dk : p->n;    g.colsB = bn ? p->n : dk;
                g.lda = an ? p->k : p->rows; g.ldb = bn ? p->n : p->k; g.ldc = p->n;
                g.alpha = alpha_host;
                g.A = an ? p->A_loc + es * (size_t)k0 : p->A_loc + es * (size_t)k0 * (size_t)p->rows;
                g.B = bn ? p->B_loc + es * (size_t)k0 * (size_t)p->n : p->B_loc + es * (size_t)k0;
                const bool last = i + 1 == nchunks;
                g.C = last ? c_root : p->C_acc;
                if (i > 0) { g.beta = one; g.Cin = p->C_acc; g.ldcin = p->n; }
                g.stream = s;
                st = gemm_device(g);
            }
        }
    }
    return st;
    };
    const int st = local_work();
    // "done": when this has run on the root's stream, every peer's last kernel -- and its stores into C -- has finished
    if (c->world > 1) ACC_NCCL(g_nccl.AllReduce(c->token, c->token, 1, NCCL_FLOAT64, NCCL_SUM, c->nccl, s));
    return st;
}

int accblas_mg_gemm(accblas_mg_comm_t comm, accblas_stream_t stream, int dtype, int opA, int opB, int64_t rowsA, int64_t colsA,
                    int64_t rowsB, int64_t colsB, const void *alpha_host, const void *A, const void *B, void *C, int root)
{
    accblas_mg_plan_t plan = nullptr;
    ACC_TRY(accblas_mg_gemm_plan(comm, dtype, opA, opB, rowsA, colsA, rowsB, colsB, A, B, C, root, &plan));
    int st = accblas_mg_gemm_run(plan, stream, alpha_host);
    const cudaError_t e = cudaStreamSynchronize(static_cast<cudaStream_t>(stream));     // the replicas die with the plan
    accblas_mg_gemm_plan_destroy(plan);
    if (st == ACCBLAS_OK && e != cudaSuccess) st = cuda_fail(e, "accblas_mg_gemm");
    return st;
}

// ------------------------------------------------------------------------------------------------------------
// gemv / dot over the exchange buffer
// ------------------------------------------------------------------------------------------------------------
int accblas_mg_gemv(accblas_mg_comm_t comm, accblas_stream_t stream, int dtype, int opA, int64_t rows_local, int64_t colsA,
                    const void *alpha_host, const void *A_local, const void *x, int64_t lenx, void *y_full)
{
    Comm *c = reinterpret_cast<Comm *>(comm);
    if (!c) return ACCBLAS_ERR_NULL;
    if (dtype < ACCBLAS_F32 || dtype > ACCBLAS_C64) return ACCBLAS_ERR_DTYPE;
    if (opA < 0 || opA > 2) return ACCBLAS_ERR_OP;
    if (rows_local < 0 || colsA < 0) return ACCBLAS_ERR_SHAPE;
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    const size_t es = elem_size(dtype);
    const bool n_op = opA == ACCBLAS_OP_N;
    const int64_t leny_local = n_op ? rows_local : colsA;
    // T / H: colsA is the same on every rank, so this return is taken by all of them or by none.  N with an empty
    // local panel still has to take part in the exchange below (the other ranks wait for this rank's flag).
    if (!n_op && leny_local == 0) return ACCBLAS_OK;
    if (!y_full && (leny_local > 0 || c->world == 1)) return ACCBLAS_ERR_NULL;
    if (es * (size_t)leny_local > XSLOT) return ACCBLAS_ERR_WORKSPACE;          // beyond the exchange slot: shard coarser
    // local product into scratch (16-byte aligned, whole vectors: the deposit is copied in 16-byte units)
    char *mine = nullptr;
    ACC_CUDA(scratch_alloc(reinterpret_cast<void **>(&mine), es * (size_t)leny_local + 16, s));
    int st = leny_local > 0 ? accblas_gemv(stream, dtype, opA, rows_local, colsA, alpha_host, A_local, x, lenx, mine) : ACCBLAS_OK;
    if (st == ACCBLAS_OK) {
        const size_t elems = (size_t)leny_local;
        if (n_op) {          // all-gather of the row slices, each rank's length travelling with its deposit
            st = dtype == ACCBLAS_F32 ? exchange<float, false>(c, s, mine, elems, static_cast<float *>(y_full))
               : dtype == ACCBLAS_F64 ? exchange<double, false>(c, s, mine, elems, static_cast<double *>(y_full))
               : dtype == ACCBLAS_C32 ? exchange<float2, false>(c, s, mine, elems, static_cast<float2 *>(y_full))
                                      : exchange<double2, false>(c, s, mine, elems, static_cast<double2 *>(y_full));
        } else {             // all-reduce of the partial y, summed in rank order
            st = dtype == ACCBLAS_F32 ? exchange<float, true>(c, s, mine, elems, static_cast<float *>(y_full))
               : dtype == ACCBLAS_F64 ? exchange<double, true>(c, s, mine, elems, static_cast<double *>(y_full))
               : dtype == ACCBLAS_C32 ? exchange<float2, true>(c, s, mine, elems, static_cast<float2 *>(y_full))
                                      : exchange<double2, true>(c, s, mine, elems, static_cast<double2 *>(y_full));
        }
    }
    const cudaError_t e = cudaFreeAsync(mine, s);
    if (st == ACCBLAS_OK && e != cudaSuccess) return cuda_fail(e, "cudaFreeAsync");
    return st;
}

int accblas_mg_dot(accblas_mg_comm_t comm, accblas_stream_t stream, int dtype, int conj_x, int64_t n_local, const void *x_local,
                   const void *y_local, void *out_dev)
{
    Comm *c = reinterpret_cast<Comm *>(comm);
    if (!c || !out_dev) return ACCBLAS_ERR_NULL;
    if (dtype < ACCBLAS_F32 || dtype > ACCBLAS_C64) return ACCBLAS_ERR_DTYPE;
    if (n_local < 0) return ACCBLAS_ERR_SHAPE;
    if (n_local > 0 && (!x_local || !y_local)) return ACCBLAS_ERR_NULL;
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    if (g_tuning[ACCBLAS_TUNE_MG_DOT_FUSED] != 0) {
        // ONE launch: the dot kernel's last CTA sums the partials and performs the exchange (dot.cu: dot_mg_kernel)
        DotExchange ex;
        ex.peers = peer_ptrs(c);
        ex.rank = c->rank; ex.world = c->world;
        ex.seq = ++c->seq;
        ex.parity_off = (ex.seq & 1u) * c->xparity;
        ex.counter = reinterpret_cast<unsigned *>(c->dot_scratch);
        ex.partials = c->dot_scratch + 64;
        return dot_mg_dispatch(s, dtype, conj_x, n_local, x_local, y_local, out_dev, ex, MG_DOT_GRID);
    }
    // round-2a path, kept for the A/B measurement: local dot (two launches) + one exchange kernel
    char *mine = nullptr;
    ACC_CUDA(scratch_alloc(reinterpret_cast<void **>(&mine), 16, s));
    int st = accblas_dot(stream, dtype, conj_x, n_local, x_local, y_local, mine);
    if (st == ACCBLAS_OK) {
        const uint32_t seq = ++c->seq;
        const size_t off = (seq & 1u) * c->xparity;
        const PeerPtrs pp = peer_ptrs(c);
        switch (dtype) {
        case ACCBLAS_F32: mg_scalar_allreduce_kernel<float><<<1, 32, 0, s>>>(pp, c->rank, c->world, off, seq, reinterpret_cast<const float *>(mine), static_cast<float *>(out_dev)); break;
        case ACCBLAS_F64: mg_scalar_allreduce_kernel<double><<<1, 32, 0, s>>>(pp, c->rank, c->world, off, seq, reinterpret_cast<const double *>(mine), static_cast<double *>(out_dev)); break;
        case ACCBLAS_C32: mg_scalar_allreduce_kernel<float2><<<1, 32, 0, s>>>(pp, c->rank, c->world, off, seq, reinterpret_cast<const float2 *>(mine), static_cast<float2 *>(out_dev)); break;
        default:          mg_scalar_allreduce_kernel<double2><<<1, 32, 0, s>>>(pp, c->rank, c->world, off, seq, reinterpret_cast<const double2 *>(mine), static_cast<double2 *>(out_dev)); break;
        }
        st = check_launch("mg_scalar_allreduce_kernel");
    }
    const cudaError_t e = cudaFreeAsync(mine, s);
    if (st == ACCBLAS_OK && e != cudaSuccess) return cuda_fail(e, "cudaFreeAsync");
    return st;
}

}  // extern "C"
