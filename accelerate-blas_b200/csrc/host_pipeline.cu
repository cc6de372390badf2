// host_pipeline.cu -- accblas_gemm_host: what `run (A <> B)` does end to end in the reference
// (`use` uploads the host arrays, the hook runs, the result is copied back:
// test/Test/BLAS/Level3.hs:52-54, bench/Bench/Accelerate.hs:111-114), as ONE pipelined call.
//
// The reference serialises upload -> cublas?gemm -> download.  Here the three overlap as a
// wavefront over a P x Q block grid of C:
//
//   step t uploads row panel t of op(A) and column panel t of op(B) (one H2D stream, so the
//   PCIe link carries exactly one copy at a time, in the order A0 B0 A1 B1 ...);
//   block (i, j) of C is computable once max(i, j) panels have landed, so step t releases an
//   L-shaped strip:  the row strip  C[t, 0..t-1]  (needs A_t, already has B_0..B_{t-1})  and
//                    the column strip C[0..t, t]   (needs B_t, already has A_0..A_t),
//   each ONE product on a sub-matrix view (row pitches lda/ldb/ldc of the full arrays), on two
//   compute streams so that partial waves of one strip are filled by the other;
//   finished strips go back over the D2H engine (full duplex with the uploads).
//
// The amount of computable work grows quadratically with the bytes landed, so the multiply is
// data-starved only during the first ~40 % of the upload and the tail after the last panel is
// 2/P of the product instead of all of it.  Panels are whole 128-multiples, so every view keeps
// the 16-byte alignment the TMA kernels want.
#include "common.cuh"
#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <string>
#include <vector>

namespace accblas {
namespace {

struct Scratch {
    void *p = nullptr;
    cudaStream_t s;
    explicit Scratch(cudaStream_t st) : s(st) {}
    cudaError_t alloc(size_t bytes) { return scratch_alloc(&p, bytes, s); }
    ~Scratch() { if (p) cudaFreeAsync(p, s); }
};

// The side streams and events of the wavefront are per-thread, per-device resources created on first
// use and kept (creating four streams costs ~0.3 ms, more than a 1000^3 product).  They carry no state
// between calls: every call starts by making them wait on the caller's stream and ends by joining them
// back, and accblas_gemm_host is synchronous, so a thread never has two pipelines in flight.
struct Streams {
    cudaStream_t up = nullptr, down = nullptr, row = nullptr, col = nullptr;
    std::vector<cudaEvent_t> events;
    size_t used = 0;
    cudaError_t create() {
        if (up) return cudaSuccess;
        cudaError_t e;
        if ((e = cudaStreamCreateWithFlags(&up, cudaStreamNonBlocking)) != cudaSuccess) return e;
        if ((e = cudaStreamCreateWithFlags(&down, cudaStreamNonBlocking)) != cudaSuccess) return e;
        if ((e = cudaStreamCreateWithFlags(&row, cudaStreamNonBlocking)) != cudaSuccess) return e;
        return cudaStreamCreateWithFlags(&col, cudaStreamNonBlocking);
    }
    cudaEvent_t make() {                       // next pooled event (valid until the next call's reset)
        if (used == events.size()) {
            cudaEvent_t e = nullptr;
            cudaEventCreateWithFlags(&e, cudaEventDisableTiming);
            events.push_back(e);
        }
        return events[used++];
    }
};

// ACCBLAS_HOST_TRACE=1: timed events at every stage boundary of the wavefront, printed (ms since the call's start) to
// stderr when the call ends -- a debugging aid for the pipeline's schedule, never on by default.
struct Trace {
    bool on;
    cudaEvent_t t0 = nullptr;
    std::vector<std::pair<std::string, cudaEvent_t>> marks;
    Trace() : on([] { const char *e = getenv("ACCBLAS_HOST_TRACE"); return e && *e == '1'; }()) {}
    void start(cudaStream_t s) { if (on) { cudaEventCreate(&t0); cudaEventRecord(t0, s); } }
    void mark(const char *what, int t, cudaStream_t s) {
        if (!on) return;
        cudaEvent_t e = nullptr;
        cudaEventCreate(&e);
        cudaEventRecord(e, s);
        marks.emplace_back(std::string(what) + std::to_string(t), e);
    }
    void dump() {
        if (!on) return;
        for (auto &m : marks) {
            float ms = 0.f;
            cudaEventElapsedTime(&ms, t0, m.second);
            fprintf(stderr, "%s=%.3f ", m.first.c_str(), ms);
            cudaEventDestroy(m.second);
        }
        fprintf(stderr, "\n");
        cudaEventDestroy(t0);
    }
};

Streams &thread_streams()
{
    static thread_local Streams per_device[64];
    int dev = 0;
    cudaGetDevice(&dev);
    Streams &s = per_device[(dev >= 0 && dev < 64) ? dev : 0];
    s.used = 0;
    return s;
}

// panel extent: a multiple of 128, at most `maxp` panels, at least 256 wide
int64_t panel_extent(int64_t extent, int maxp)
{
    int64_t w = std::max<int64_t>(256, (extent + maxp - 1) / maxp);
    return (w + 127) / 128 * 128;
}

// copy a rows x cols window (element offsets r0, c0) of a tight row-major rows_total x ld matrix
cudaError_t copy_window(void *dst, const void *src, size_t es, int64_t ld, int64_t r0, int64_t c0,
                        int64_t rows, int64_t cols, cudaMemcpyKind kind, cudaStream_t s)
{
    if (rows <= 0 || cols <= 0) return cudaSuccess;
    const size_t off = es * ((size_t)r0 * (size_t)ld + (size_t)c0);
    char *d = static_cast<char *>(dst) + off;
    const char *p = static_cast<const char *>(src) + off;
    if (cols == ld) return cudaMemcpyAsync(d, p, es * (size_t)rows * (size_t)ld, kind, s);
    return cudaMemcpy2DAsync(d, es * (size_t)ld, p, es * (size_t)ld, es * (size_t)cols, (size_t)rows, kind, s);
}

}  // namespace

int gemm_host_pipeline(const GemmArgs &g0, const void *A_host, const void *B_host, void *C_host)
{
    const GemmArgs &g = g0;                      // shapes derived, H on real types already folded to T
    const size_t es = elem_size(g.dtype);
    cudaStream_t s = g.stream;
    const size_t bytesA = es * (size_t)g.rowsA * (size_t)g.colsA, bytesB = es * (size_t)g.rowsB * (size_t)g.colsB;
    const size_t bytesC = es * (size_t)g.m * (size_t)g.n;
    Scratch dA(s), dB(s), dC(s);
    ACC_CUDA(dA.alloc(bytesA));
    ACC_CUDA(dB.alloc(bytesB));
    ACC_CUDA(dC.alloc(bytesC));

    // the tcgen05 family re-packs the operand views of every strip: fewer, larger panels there
    const int maxp = g_tuning[ACCBLAS_TUNE_HOST_PANELS] > 0 ? g_tuning[ACCBLAS_TUNE_HOST_PANELS]
                   : (g.dtype == ACCBLAS_F32 || g.dtype == ACCBLAS_C32) ? 8 : 32;   // swept: tools/exp/host_panels.py
    const int64_t pm = panel_extent(g.m, maxp), pn = panel_extent(g.n, maxp);
    const int P = (int)((g.m + pm - 1) / pm), Q = (int)((g.n + pn - 1) / pn);

    // below ~1 ms of PCIe traffic the wavefront's extra launches and events cost more than they hide
    const int force = g_tuning[ACCBLAS_TUNE_HOST_PIPELINE];
    const bool small = force == 1 ? false : force == 0 ? true : bytesA + bytesB + bytesC < ((size_t)48 << 20);
    if (P * Q == 1 || g.k == 0 || small) {
        if (bytesA) ACC_CUDA(cudaMemcpyAsync(dA.p, A_host, bytesA, cudaMemcpyHostToDevice, s));
        if (bytesB) ACC_CUDA(cudaMemcpyAsync(dB.p, B_host, bytesB, cudaMemcpyHostToDevice, s));
        GemmArgs one = g;
        one.A = dA.p; one.B = dB.p; one.C = dC.p;
        ACC_TRY(gemm_device(one));
        ACC_CUDA(cudaMemcpyAsync(C_host, dC.p, bytesC, cudaMemcpyDeviceToHost, s));
        ACC_CUDA(cudaStreamSynchronize(s));
        return ACCBLAS_OK;
    }

    Streams &st = thread_streams();
    ACC_CUDA(st.create());
    Streams &evs = st;
    cudaEvent_t ev_start = evs.make();
    Trace trace;
    trace.start(s);
    ACC_CUDA(cudaEventRecord(ev_start, s));      // the allocations are ordered on s
    for (cudaStream_t x : {st.up, st.down, st.row, st.col}) ACC_CUDA(cudaStreamWaitEvent(x, ev_start, 0));

    const bool an = g.opA == ACCBLAS_OP_N, bn = g.opB == ACCBLAS_OP_N;
    // sub-product on views: op(A) rows [r0, r0+h), op(B) columns [c0, c0+w)
    auto product = [&](int64_t r0, int64_t h, int64_t c0, int64_t w, cudaStream_t on) -> int {
        GemmArgs v = g;
        v.stream = on;
        v.no_persistent = true;                  // two strips multiply concurrently: a one-CTA-per-SM schedule would serialise them
        v.lda = g.colsA; v.ldb = g.colsB; v.ldc = g.n;
        if (an) { v.A = static_cast<char *>(dA.p) + es * (size_t)r0 * (size_t)g.colsA; v.rowsA = h; }
        else    { v.A = static_cast<char *>(dA.p) + es * (size_t)r0; v.colsA = h; }
        if (bn) { v.B = static_cast<char *>(dB.p) + es * (size_t)c0; v.colsB = w; }
        else    { v.B = static_cast<char *>(dB.p) + es * (size_t)c0 * (size_t)g.colsB; v.rowsB = w; }
        v.C = static_cast<char *>(dC.p) + es * ((size_t)r0 * (size_t)g.n + (size_t)c0);
        return gemm_device(v);
    };

    int status = ACCBLAS_OK;
    cudaError_t ce = cudaSuccess;
    auto ck = [&](cudaError_t e) { if (ce == cudaSuccess && e != cudaSuccess) ce = e; };
    const int steps = std::max(P, Q);
    for (int t = 0; t < steps && status == ACCBLAS_OK && ce == cudaSuccess; ++t) {
        const int64_t r0 = (int64_t)t * pm, h = t < P ? std::min(pm, g.m - r0) : 0;
        const int64_t c0 = (int64_t)t * pn, w = t < Q ? std::min(pn, g.n - c0) : 0;
        cudaEvent_t ev_a = nullptr, ev_b = nullptr;
        if (h > 0) {                              // row panel t of op(A)
            ck(an ? copy_window(dA.p, A_host, es, g.colsA, r0, 0, h, g.colsA, cudaMemcpyHostToDevice, st.up)
                  : copy_window(dA.p, A_host, es, g.colsA, 0, r0, g.rowsA, h, cudaMemcpyHostToDevice, st.up));
            ev_a = evs.make();
            ck(cudaEventRecord(ev_a, st.up));
            trace.mark("upA", t, st.up);
        }
        if (w > 0) {                              // column panel t of op(B)
            ck(bn ? copy_window(dB.p, B_host, es, g.colsB, 0, c0, g.rowsB, w, cudaMemcpyHostToDevice, st.up)
                  : copy_window(dB.p, B_host, es, g.colsB, c0, 0, w, g.colsB, cudaMemcpyHostToDevice, st.up));
            ev_b = evs.make();
            ck(cudaEventRecord(ev_b, st.up));
            trace.mark("upB", t, st.up);
        }
        // row strip C[t, 0 .. min(t,Q)-1]: needs A_t (B_0..B_{t-1} precede it on the upload stream)
        const int64_t wn = std::min<int64_t>((int64_t)std::min(t, Q) * pn, g.n);
        if (h > 0 && wn > 0) {
            ck(cudaStreamWaitEvent(st.row, ev_a, 0));
            status = product(r0, h, 0, wn, st.row);
            cudaEvent_t done = evs.make();
            ck(cudaEventRecord(done, st.row));
            trace.mark("row", t, st.row);
            ck(cudaStreamWaitEvent(st.down, done, 0));
            ck(copy_window(C_host, dC.p, es, g.n, r0, 0, h, wn, cudaMemcpyDeviceToHost, st.down));
        }
        // column strip C[0 .. min(t+1,P)-1, t]: needs B_t (A_0..A_t precede it)
        const int64_t hm = std::min<int64_t>((int64_t)std::min(t + 1, P) * pm, g.m);
        if (w > 0 && hm > 0 && status == ACCBLAS_OK) {
            ck(cudaStreamWaitEvent(st.col, ev_b, 0));
            status = product(0, hm, c0, w, st.col);
            cudaEvent_t done = evs.make();
            ck(cudaEventRecord(done, st.col));
            trace.mark("col", t, st.col);
            ck(cudaStreamWaitEvent(st.down, done, 0));
            ck(copy_window(C_host, dC.p, es, g.n, 0, c0, hm, w, cudaMemcpyDeviceToHost, st.down));
        }
    }
    trace.mark("down", steps, st.down);
    // join everything back into the caller's stream (the scratch is freed on it), then wait
    for (cudaStream_t x : {st.up, st.row, st.col, st.down}) {
        cudaEvent_t e = evs.make();
        ck(cudaEventRecord(e, x));
        ck(cudaStreamWaitEvent(s, e, 0));
    }
    ck(cudaStreamSynchronize(s));
    for (cudaStream_t x : {st.up, st.row, st.col, st.down}) cudaStreamSynchronize(x);
    trace.dump();
    if (status != ACCBLAS_OK) return status;
    if (ce != cudaSuccess) return cuda_fail(ce, "accblas_gemm_host pipeline");
    return ACCBLAS_OK;
}

}  // namespace accblas
