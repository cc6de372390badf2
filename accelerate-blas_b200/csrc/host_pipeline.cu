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
//
// COMPUTE-BOUND products (round 2; dgemm 8192^3: 30.5 ms of multiply against 19.3 ms of upload) take a second schedule.
// The wavefront starves the GPU early -- strip t holds 2t + 1 blocks, each a full-K chain: the first 18 strips are one
// partial wave each -- and measured 36.2 ms where upload-then-overlap arithmetic allows ~32.  Instead:
//
//   phase 1 ("rank updates")   K is cut into panels of growing width (256, 256, 512, 1024, ...).  Panel p uploads the K-slab
//                              of op(A) and of op(B); as soon as it has landed the WHOLE C receives
//                              C (+)= op(A)[:, panel] . op(B)[panel, :]  (accblas_gemm_ex, beta = 1 in place after the
//                              first panel) -- a full grid of tiles from the first 0.6 ms on.  Multiplication is the
//                              slower side, so it never waits again; phase 1 ends after the fraction of K whose products
//                              take as long as the whole upload (~63 % for dgemm 8192^3).
//   phase 2 ("row finish")     everything is on the device.  The remaining K range is applied row panel by row panel,
//                              C[panel, :] += op(A)[panel, rest] . op(B)[rest, :], and every finished panel goes back over
//                              the D2H engine while the next one multiplies; only the last panel's download is exposed.
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


// nominal rates for choosing the schedule (not a performance claim: only their ratio to the link rate matters)
double nominal_flops_per_s(int dtype)
{
    return dtype == ACCBLAS_F64 ? 35e12 : dtype == ACCBLAS_C64 ? 44e12 : dtype == ACCBLAS_F32 ? 230e12 : 270e12;
}
constexpr double kLinkBytesPerS = 55e9;          // pinned-memory PCIe rate of the B200 boxes (measured 55.5 GB/s H2D, 57 D2H)

struct KPanelPlan {
    bool use = false;
    std::vector<int64_t> kcut;                   // phase-1 panel boundaries: kcut[0] = 0 < ... < kcut.back() = k1
    int64_t k1 = 0;                              // phase 2 multiplies K range [k1, k)
    int64_t row_panel = 0;                       // rows of op(A) / C per phase-2 product
};

KPanelPlan plan_kpanels(const GemmArgs &g, size_t bytesA, size_t bytesB, size_t bytesC)
{
    KPanelPlan pl;
    const int knob = g_tuning[ACCBLAS_TUNE_HOST_SCHEDULE];
    if (knob == 0 || g.k < 2048 || g.m < 1024 || g.n < 1024) return pl;
    const double flops = (is_complex(g.dtype) ? 8.0 : 2.0) * (double)g.m * (double)g.n * (double)g.k;
    const double t_comp = flops / nominal_flops_per_s(g.dtype), t_up = (double)(bytesA + bytesB) / kLinkBytesPerS,
                 t_down = (double)bytesC / kLinkBytesPerS;
    if (knob != 1 && t_comp < 1.3 * t_up) return pl;          // transfer-bound: the wavefront overlaps up- and download
    // phase 1 lasts as long as the upload; phase 2 must also be long enough to hide the download of all but one panel
    double f = std::min(t_up / t_comp * 1.03, 1.0 - std::min(0.9, t_down / t_comp));
    f = std::max(0.25, std::min(0.8, f));
    const int64_t k1_target = (int64_t)(f * (double)g.k);
    // widths 256, 256, 512, 1024, 1024, ...: narrow first so that multiplication starts 0.6 ms into the call, wider later
    // because every pass re-reads and re-writes all of C (~0.3 ms each) -- but not wider than 1024: multiplication is only
    // ~1 ms behind the link in the first third of the call, and a 2048-wide slab (4.8 ms of upload) made it wait 0.75 ms.
    // The last panel is clipped to the target.
    const int64_t k1_end = std::max<int64_t>(256, (k1_target + 128) / 256 * 256);
    int64_t w = 256, at = 0;
    pl.kcut.push_back(0);
    int same = 0;
    while (at < k1_end) {
        int64_t next = std::min<int64_t>(at + w, k1_end);
        if (k1_end - next < 256) next = k1_end;                 // no sliver at the end
        if (g.k - next < 256) next = g.k;
        pl.kcut.push_back(next);
        at = next;
        if (++same == 2 || w >= 512) { same = 0; if (w < 1024) w *= 2; }
    }
    pl.k1 = at;
    if (pl.k1 >= g.k) { pl.kcut.clear(); return pl; }          // nothing left for phase 2: keep the wavefront
    pl.row_panel = panel_extent(g.m, 16);
    if (const char *e = getenv("ACCBLAS_HOST_ROWPANEL")) { const int64_t v = atoll(e); if (v >= 128) pl.row_panel = v / 128 * 128; }   // experiments
    pl.use = true;
    return pl;
}

int gemm_host_kpanels(const GemmArgs &g, const void *A_host, const void *B_host, void *C_host, void *dA, void *dB, void *dC,
                      const KPanelPlan &pl)
{
    const size_t es = elem_size(g.dtype);
    cudaStream_t s = g.stream;
    Streams &st = thread_streams();
    ACC_CUDA(st.create());
    cudaEvent_t ev_start = st.make();
    Trace trace;
    trace.start(s);
    ACC_CUDA(cudaEventRecord(ev_start, s));
    cudaStream_t up = st.up, down = st.down, comp = st.row;
    for (cudaStream_t x : {up, down, comp}) ACC_CUDA(cudaStreamWaitEvent(x, ev_start, 0));
    const bool an = g.opA == ACCBLAS_OP_N, bn = g.opB == ACCBLAS_OP_N;
    static const double one_d[2] = {1.0, 0.0};
    static const float one_f[2] = {1.f, 0.f};
    const void *one = (g.dtype == ACCBLAS_F32 || g.dtype == ACCBLAS_C32) ? static_cast<const void *>(one_f) : static_cast<const void *>(one_d);
    cudaError_t ce = cudaSuccess;
    auto ck = [&](cudaError_t e) { if (ce == cudaSuccess && e != cudaSuccess) ce = e; };
    int status = ACCBLAS_OK;

    // op(A)[r0 : r0 + h, k0 : k1] . op(B)[k0 : k1, :] added into (or, first = true, written to) C[r0 : r0 + h, :]
    auto product = [&](int64_t r0, int64_t h, int64_t k0, int64_t k1, bool first, cudaStream_t on) -> int {
        GemmArgs v = g;
        v.stream = on;
        v.no_persistent = on != comp || !first;  // (beta products take the data-parallel tiles anyway; phase 2 runs two at a time)
        v.lda = g.colsA; v.ldb = g.colsB; v.ldc = g.n;
        const int64_t w = k1 - k0;
        if (an) { v.A = static_cast<char *>(dA) + es * ((size_t)r0 * (size_t)g.colsA + (size_t)k0); v.rowsA = h; v.colsA = w; }
        else    { v.A = static_cast<char *>(dA) + es * ((size_t)k0 * (size_t)g.colsA + (size_t)r0); v.rowsA = w; v.colsA = h; }
        if (bn) { v.B = static_cast<char *>(dB) + es * (size_t)k0 * (size_t)g.colsB; v.rowsB = w; }
        else    { v.B = static_cast<char *>(dB) + es * (size_t)k0; v.colsB = w; }
        v.C = static_cast<char *>(dC) + es * (size_t)r0 * (size_t)g.n;
        if (!first) { v.beta = one; v.Cin = v.C; v.ldcin = g.n; }
        return gemm_device(v);
    };
    // upload the K-slab [k0, k1) of both operands; returns the event that marks it landed
    auto upload = [&](int64_t k0, int64_t k1, int tag) -> cudaEvent_t {
        const int64_t w = k1 - k0;
        ck(an ? copy_window(dA, A_host, es, g.colsA, 0, k0, g.rowsA, w, cudaMemcpyHostToDevice, up)
              : copy_window(dA, A_host, es, g.colsA, k0, 0, w, g.colsA, cudaMemcpyHostToDevice, up));
        ck(bn ? copy_window(dB, B_host, es, g.colsB, k0, 0, w, g.colsB, cudaMemcpyHostToDevice, up)
              : copy_window(dB, B_host, es, g.colsB, 0, k0, g.rowsB, w, cudaMemcpyHostToDevice, up));
        cudaEvent_t e = st.make();
        ck(cudaEventRecord(e, up));
        trace.mark("up", tag, up);
        return e;
    };

    // ---- phase 1: rank updates of the whole C, panel by panel as the slabs land ----
    // f64: on the 64 x 128 tile at two CTAs per SM -- one CTA's beta epilogue (read C, add, write C: ~8 us of every ~140 us
    // tile at K = 1024) runs under the other's main loop: 4.10 ms per 1024-wide pass against 4.21 on the 128 x 128 tile
    // (ideal 3.81); phase 2's long-K products keep the automatic choice (128 x 128: 0.78 ms per row panel against 0.90).
    const int P1 = (int)pl.kcut.size() - 1;
    const int saved_tile = g_tuning[ACCBLAS_TUNE_DMMA_TILE];
    if (g.dtype == ACCBLAS_F64 && saved_tile < 0) g_tuning[ACCBLAS_TUNE_DMMA_TILE] = 1;
    for (int p = 0; p < P1 && status == ACCBLAS_OK && ce == cudaSuccess; ++p) {
        cudaEvent_t landed = upload(pl.kcut[p], pl.kcut[p + 1], p);
        ck(cudaStreamWaitEvent(comp, landed, 0));
        status = product(0, g.m, pl.kcut[p], pl.kcut[p + 1], p == 0, comp);
        trace.mark("upd", p, comp);
    }
    g_tuning[ACCBLAS_TUNE_DMMA_TILE] = saved_tile;
    // ---- the rest of both operands (needed in full by every phase-2 product), in a few pieces so that the copies
    //      queue behind the phase-1 slabs in K order ----
    cudaEvent_t all_landed = nullptr;
    for (int64_t k0 = pl.k1, tag = P1; k0 < g.k && ce == cudaSuccess; ++tag) {
        const int64_t k1 = std::min<int64_t>(g.k, k0 + 2048);
        all_landed = upload(k0, k1, (int)tag);
        k0 = k1;
    }
    // ---- phase 2: finish C row panel by row panel, download each as soon as it is done ----
    // The panels are independent products of 256 tiles each (1.73 waves of 148): they alternate between two streams so
    // that the next panel's CTAs fill the SMs the partial second wave leaves idle.
    cudaStream_t comp2 = st.col;
    ck(cudaStreamWaitEvent(comp2, ev_start, 0));
    {
        cudaEvent_t p1_done = st.make();
        ck(cudaEventRecord(p1_done, comp));
        ck(cudaStreamWaitEvent(comp2, p1_done, 0));
    }
    if (all_landed) { ck(cudaStreamWaitEvent(comp, all_landed, 0)); ck(cudaStreamWaitEvent(comp2, all_landed, 0)); }
    int idx = 0;
    for (int64_t r0 = 0; r0 < g.m && status == ACCBLAS_OK && ce == cudaSuccess; r0 += pl.row_panel, ++idx) {
        const int64_t h = std::min<int64_t>(pl.row_panel, g.m - r0);
        cudaStream_t on = (idx & 1) ? comp2 : comp;
        status = product(r0, h, pl.k1, g.k, false, on);
        cudaEvent_t done = st.make();
        ck(cudaEventRecord(done, on));
        trace.mark("fin", idx, on);
        ck(cudaStreamWaitEvent(down, done, 0));
        ck(copy_window(C_host, dC, es, g.n, r0, 0, h, g.n, cudaMemcpyDeviceToHost, down));
    }
    trace.mark("down", idx, down);
    for (cudaStream_t x : {up, comp, comp2, down}) {
        cudaEvent_t e = st.make();
        ck(cudaEventRecord(e, x));
        ck(cudaStreamWaitEvent(s, e, 0));
    }
    ck(cudaStreamSynchronize(s));
    for (cudaStream_t x : {up, comp, comp2, down}) cudaStreamSynchronize(x);
    trace.dump();
    if (status != ACCBLAS_OK) return status;
    if (ce != cudaSuccess) return cuda_fail(ce, "accblas_gemm_host K-panel pipeline");
    return ACCBLAS_OK;
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

    {
        const KPanelPlan kp = plan_kpanels(g, bytesA, bytesB, bytesC);
        if (kp.use) return gemm_host_kpanels(g, A_host, B_host, C_host, dA.p, dB.p, dC.p, kp);
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
