// gemm_dmma.cu -- FP64 GEMM on the DMMA tensor path (mma.sync m8n8k4 f64 -> SASS DMMA.8x8x4),
// TMA-staged, warp-specialised.  Replaces BLAS.dgemm / BLAS.zgemm at LLVM/PTX/Level3.hs:96,106.
//
// tcgen05 has no f64 kind, so FP64 tensor work on sm_100a is the warp-level DMMA; measured
// issue-bound peak on B200 is 36.9 TFLOP/s (baseline/ubench.cu), the same 64 FMA/clk/SM the
// DFMA pipe has -- the tensor form wins by needing 1 issue slot and 2 operand registers per
// 256 FMAs, which leaves the LSU and issue ports free for operand delivery.
//
// Shape of the kernel (real f64; complex c64 further below):
//   * CTA tile 128 x 128, K-block 16 (128 bytes of a row), 6-stage ring of 32 KiB stages.
//   * warp 8 = producer: one elected lane waits `empty[s]`, arms `full[s]` with the byte count and
//     issues the TMA boxes (cp.async.bulk.tensor.2d, 128-byte swizzle, zero fill outside the matrix
//     -- so ragged m, n, k need no special code on the load side).
//   * warps 0-7 = consumers, 2 (m) x 4 (n), warp tile 64 x 32 = 8 x 4 DMMA tiles, 64 accumulator
//     doubles per lane.  Fragments come straight from the swizzled stage with ld.shared.f64.
//   * Operands are used in their STORED layout, whatever the op:
//       K-major  (A with op N, B with op T): box 16 (k) x 128 rows, one TMA per stage;
//       MN-major (A with op T, B with op N): boxes 16 (m|n) x 16 (k), eight TMAs per stage.
//   * The four k indices one DMMA consumes are NOT consecutive: lane slot t of step s reads
//       k(s,0) = 2s, k(s,1) = 2((s+1)&3)+1, k(s,2) = 8+2((s+2)&3), k(s,3) = 8+2((s+3)&3)+1
//     (the same for both operands, so the product is unchanged).  With the 128-byte swizzle this
//     set makes every half-warp ld.shared.f64 hit 16 distinct 8-byte bank pairs for BOTH majors.
//   * Epilogue: alpha * acc written with 16-byte stores straight from the fragment layout
//     (C is write-only: beta = 0, LLVM/PTX/Level3.hs:90).
#include "common.cuh"
#include "tma.cuh"
#include <algorithm>

namespace accblas {
namespace {

constexpr int BK = 16;
// CTA tiles.  A tile is WM x WN consumer warps, each owning MI x NJ DMMA tiles (8 x 8):
//   128 x 128 (2x4 warps of 64 x 32): the throughput shape, one CTA per SM, 6 stages;
//    64 x 128 (2x4 warps of 32 x 32): two CTAs per SM, for products whose 128 x 128 tiling leaves SMs
//             idle in the last wave (dgemm 1024^3 is 64 big tiles on 148 SMs);
//    64 x  64 (2x4 warps of 32 x 16) and 32 x 32 (2x2 warps of 16 x 16): the FP64 pipe runs 64 FMA/clk/SM,
//             so even a 100^3 product is compute-bound on the two SMs a 128-wide tiling would use --
//             small products are spread over as many SMs as possible (the reference's own benchmark
//             shapes are 100^3 .. 1000^3, bench/Bench/Accelerate.hs:92-116).
template <int MI_, int NJ_, int WM_, int WN_, int STAGES_, int MIN_CTAS_> struct Tile {
    static constexpr int MI = MI_, NJ = NJ_, WM = WM_, WN = WN_;
    static constexpr int BM = 8 * MI * WM, BN = 8 * NJ * WN;
    static constexpr int CONSUMER_WARPS = WM * WN;
    static constexpr int THREADS = (CONSUMER_WARPS + 1) * 32;
    static constexpr int STAGES = STAGES_;
    static constexpr int A_BYTES = BM * BK * 8, B_BYTES = BN * BK * 8;
    static constexpr int STAGE_BYTES = A_BYTES + B_BYTES;
    static constexpr int SMEM_BYTES = STAGES * STAGE_BYTES + 1024 /*align slack*/ + 2 * STAGES * 8;
    static constexpr int MIN_CTAS = MIN_CTAS_;
    static_assert(MI % 2 == 0 && NJ % 2 == 0, "MN-major operands come in 16-wide boxes");
};
using Tile128x128 = Tile<8, 4, 2, 4, 6, 1>;
using Tile64x128  = Tile<4, 4, 2, 4, 4, 2>;
using Tile64x64   = Tile<4, 2, 2, 4, 4, 3>;
using Tile32x32   = Tile<2, 2, 2, 2, 3, 7>;   // 7 CTAs per SM: 1024 tiles (dgemm 1024^3) are ONE resident wave
// The same tile with an 8-deep ring for grids of at most three CTAs per SM (500^3: 256 tiles).  There a CTA's progress is
// bound by the LATENCY of its TMA ring, not by the pipe: three 8 KiB stages in flight per CTA are ~3 K-blocks per L2
// round trip (r01_dgemm_nn_500_ncu.txt: DMMA 60 % active while 256 CTAs sit on 148 SMs, 16.3 us for 6.7 us of math).
using Tile32x32Deep = Tile<2, 2, 2, 2, 8, 3>;
// (Round 2, measured and NOT kept: dealing the K-blocks to two groups of consumer warps inside the CTA -- 250^3: 7.10 -> 6.91 us
//  of kernel time, 250 x 250 x 1024: 18.9 -> 18.8; splitting the accumulators by k-step -- 1000^3: 70 -> 82 us; loading the
//  panels with plain loads instead of the ring -- no change; a 32 x 16 tile (twice the CTAs, half the work each) -- 7.04 us;
//  polling the barriers with test_wait instead of try_wait -- 7.17; three TMA boxes per K-block instead of two -- 7.10.
//  The one-wave kernel costs ~295 ns per K-block whatever is changed.  profiles/r02_small_shapes.md.)

__device__ __forceinline__ void dmma(double &d0, double &d1, double a, double b) {
    asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};"
                 : "+d"(d0), "+d"(d1) : "d"(a), "d"(b));
}
__device__ __forceinline__ double lds64(uint32_t addr) {
    double v;
    asm volatile("ld.shared.f64 %0, [%1];" : "=d"(v) : "r"(addr));
    return v;
}
// k index that lane slot t reads in step s (see header)
__device__ __forceinline__ int kslot(int s, int t) {
    return (t & 2) * 4 + 2 * ((s + t) & 3) + (t & 1);
}

// A consumer warp hands a stage back to the TMA producer with ONE mbarrier arrive by lane 0.  ptxas hoists that
// arrive to right behind the last ld.shared of the K-block -- above the mma.sync instructions that consume the
// loaded registers -- so at the moment the producer sees it, fragment loads of the stage can still be in flight (or,
// for lanes that left the try_wait spin loop late, not even issued), and its next TMA write into the stage races
// with them.  Found in round 1 with many small CTAs per SM (tests/test_parity_gpu.py::
// test_gemm_many_small_tiles_short_k); three ways to close it (REL, selectable with ACCBLAS_TUNE_RELEASE):
//   REL_FENCED  (0): fence.proxy.async + __syncwarp + arrive.  The fence drains the warp's loads.  Costs ~0.5 % of
//                    the 128 x 128 tile's time (measured round 2: 30.74 -> 30.59 ms at 8192^3 without it).
//   REL_LIGHT   (1): __syncwarp + arrive.  NOT SOUND, not instantiated by the dispatcher: 3 mismatches in 720 runs of the
//                    32 x 32 complex tile and a run-to-run difference of the persistent kernel at 2048^3 (round 2,
//                    profiles/r02_dgemm_release_schedule.jsonl) -- the loads CAN still be in flight.  It is only the
//                    second half of the delayed release below.
//   REL_DELAYED (2): the stage of K-block i is released at the top of K-block i + 1, after the wait on the next
//                    stage's full barrier.  A warp issues in order and an mma.sync cannot issue before its operand
//                    registers are written, so by then every ld.shared of stage i has completed: no fence needed, sound
//                    by construction, for one stage less of prefetch depth.  The release sits in another basic block
//                    (behind the spin loop), so ptxas cannot hoist it above the MMAs.
enum { REL_FENCED = 0, REL_LIGHT = 1, REL_DELAYED = 2 };
template <int REL> __device__ __forceinline__ void release_stage(uint32_t empty_bar, int lane) {
    if (REL == REL_FENCED) tma::fence_proxy_async();
    __syncwarp();
    if (lane == 0) tma::mbar_arrive(empty_bar);
}
// the two halves of a K-block's release protocol.  Under REL_DELAYED the stage released at the top of a K-block is the
// PREVIOUS ring slot (derived from the current one: no register is carried across the loop body, which has none to spare).
template <int REL, int NSTAGES> __device__ __forceinline__ void release_before(uint32_t empty_base, int s, bool first, int lane) {
    if (REL == REL_DELAYED && !first) release_stage<REL_LIGHT>(empty_base + 8 * (s == 0 ? NSTAGES - 1 : s - 1), lane);
}
template <int REL> __device__ __forceinline__ void release_after(uint32_t empty_bar, int lane) {
    if (REL != REL_DELAYED) release_stage<REL>(empty_bar, lane);
}

struct DgemmParams {
    int64_t m, n, k, ldc;
    int a_one_box, b_one_box;   // MN-major operand fetched as ONE 3-D TMA box per stage (tma::make_mn3d)
    // strided batch (accblas_gemm_strided_batched): gridDim.x = batch * tiles_per_batch, the tensor maps are 3-D
    // (tma::make_2d_batched) and product b writes C + b * stride_c.  tiles_per_batch == 0: one product, 2-D maps.
    int tiles_per_batch, plane_a, plane_b;   // plane_x = 1: operand x advances with the batch index (0: shared)
    int64_t stride_c;
    // split-K (one product, 2-D maps): gridDim.x = slices * tiles_per_batch, slice s multiplies K-blocks
    // [s * kb_per_slice, (s + 1) * kb_per_slice) and writes plane s of the scratch (C + s * stride_c).  0: whole K.
    int kb_per_slice;
    Alpha<double> alpha;
    BetaC<double> beta;  // + beta * Cin (accblas_gemm_ex); cin == nullptr: C is write-only
    double *C;
    int tiles_m, tiles_n, group;
    int snake_waves;     // CTAs per resident wave (0: every tile walks K forwards)
    const double *Ag, *Bg;   // DIRECT kernels only: the stored operands and their row pitches (elements)
    int64_t lda, ldb;
};
constexpr int kDirectThreads = 512;      // DIRECT kernels: 16 loader warps, the first CONSUMER_WARPS of them go on to multiply
constexpr int kDirectMaxKBlocks = 24;    // 24 stages of 8 KiB + slack fit a CTA's 227 KiB (k <= 384 doubles)

// tile index -> (tile_m, tile_n) under the grouped rasterisation: p.group tile-rows per group, so that one resident
// wave re-uses its A and B panels out of L2
__device__ __forceinline__ void tile_coords(const DgemmParams &p, int tile, int &tile_m, int &tile_n) {
    const int per_group = p.group * p.tiles_n;
    const int grp = tile / per_group;
    const int first_m = grp * p.group;
    const int gsz = min(p.tiles_m - first_m, p.group);
    const int in_grp = tile - grp * per_group;
    tile_m = first_m + in_grp % gsz;
    tile_n = in_grp / gsz;
}

// C = alpha * acc (+ beta * Cin) from the DMMA fragment layout, 16-byte stores where the pitch allows
// BETA is a TEMPLATE parameter of the data-parallel kernels: with the beta term behind a run-time test in the one
// epilogue, ptxas gave the small tiles (56-register cap at 7 CTAs / SM) a K loop that recomputes its shared-memory
// addresses every K-block -- 141 instead of 82 instructions per 16 DMMAs in the 32 x 32 tile, 11 % on every strip of the
// host wavefront and on 1024^3 (round 2, profiles/r02_strip_ab.jsonl).
template <typename TL, bool BETA = true, int IB = 2>
__device__ __forceinline__ void dgemm_epilogue(const DgemmParams &p, const double (&acc)[TL::MI][TL::NJ][2],
                                               int m0, int n0, int wm, int wn, int g, int t, int64_t c_off = 0)
{
    constexpr int MI = TL::MI, NJ = TL::NJ;
    const bool vec = ((p.n | p.ldc) & 1) == 0;
    const double alpha = p.alpha.get();
    const bool acc_in = BETA && p.beta.on();
    double *const Cb = p.C + c_off;
    if constexpr (IB > 0) if (acc_in) {
        // beta * Cin: Cin may be C itself, so the compiler must keep every load behind the preceding store -- one
        // HBM / L2 round trip per 16 bytes, 32 of them in a row for a 128 x 128 tile (measured: ~20 us per tile, 0.55 ms per
        // in-place pass over an 8192^2 C; the K-panel host schedule and the multi-GPU chunk pipeline make many such
        // passes).  A thread only ever reads the elements it writes itself, so the loads of IB rows are issued by hand
        // before the first store of the batch: IB * NJ round trips overlap.  (IB = 1 in the persistent kernel, whose main loop
        // sits at the 168-register limit of a 9-warp CTA: a deeper batch there spilt in the fenced variants too.)
        static_assert(IB == 0 || MI % (IB > 0 ? IB : 1) == 0, "row batches divide the warp tile");
        const bool cvec = vec && ((p.beta.ld & 1) == 0) && ((reinterpret_cast<uintptr_t>(p.beta.cin) & 15u) == 0);
        const double *const Cin = p.beta.cin;          // (no batch offset: the strided batch has no beta term)
        const double beta = p.beta.beta;
#pragma unroll
        for (int i0 = 0; i0 < MI; i0 += IB) {
            double cin[IB][NJ][2];
#pragma unroll
            for (int ii = 0; ii < IB; ++ii) {
                const int64_t r = m0 + wm * MI * 8 + (i0 + ii) * 8 + g;
#pragma unroll
                for (int j = 0; j < NJ; ++j) {
                    const int64_t c = n0 + wn * NJ * 8 + j * 8 + 2 * t;
                    cin[ii][j][0] = cin[ii][j][1] = 0.0;
                    if (r < p.m) {
                        const double *src = Cin + r * p.beta.ld + c;
                        if (cvec) {
                            if (c < p.n) { const double2 v = *reinterpret_cast<const double2 *>(src); cin[ii][j][0] = v.x; cin[ii][j][1] = v.y; }
                        } else {
                            if (c < p.n) cin[ii][j][0] = src[0];
                            if (c + 1 < p.n) cin[ii][j][1] = src[1];
                        }
                    }
                }
            }
#pragma unroll
            for (int ii = 0; ii < IB; ++ii) {
                const int i = i0 + ii;
                const int64_t r = m0 + wm * MI * 8 + i * 8 + g;
                if (r >= p.m) continue;
                double *crow = Cb + r * p.ldc;
#pragma unroll
                for (int j = 0; j < NJ; ++j) {
                    const int64_t c = n0 + wn * NJ * 8 + j * 8 + 2 * t;
                    const double v0 = fma(beta, cin[ii][j][0], alpha * acc[i][j][0]);
                    const double v1 = fma(beta, cin[ii][j][1], alpha * acc[i][j][1]);
                    if (vec) {
                        if (c < p.n) *reinterpret_cast<double2 *>(crow + c) = make_double2(v0, v1);
                    } else {
                        if (c < p.n) crow[c] = v0;
                        if (c + 1 < p.n) crow[c + 1] = v1;
                    }
                }
            }
        }
        return;
    }
#pragma unroll
    for (int i = 0; i < MI; ++i) {
        const int64_t r = m0 + wm * MI * 8 + i * 8 + g;
        if (r >= p.m) continue;
        double *crow = Cb + r * p.ldc;
#pragma unroll
        for (int j = 0; j < NJ; ++j) {
            const int64_t c = n0 + wn * NJ * 8 + j * 8 + 2 * t;
            double v0 = alpha * acc[i][j][0], v1 = alpha * acc[i][j][1];
            if (IB == 0 && acc_in) {              // one element at a time (the persistent kernel)
                if (c < p.n) v0 = p.beta.add(v0, r, c);
                if (c + 1 < p.n) v1 = p.beta.add(v1, r, c + 1);
            }
            if (vec) {
                if (c < p.n) *reinterpret_cast<double2 *>(crow + c) = make_double2(v0, v1);
            } else {
                if (c < p.n) crow[c] = v0;
                if (c + 1 < p.n) crow[c + 1] = v1;
            }
        }
    }
}

// AK: op(A) is K-major in memory (opA == N).  BKM: op(B) is K-major in memory (opB == T).
// DIRECT (round 2): operands that miss TMA's 16-byte rule (odd column counts: 8-byte aligned rows) in a product of at most
// one 32 x 32 tile per SM with k <= 384 -- no TMA ring, no re-pitch copy in front (api.cu: gemm_repitched).  All 16 warps
// of the CTA load its two panels for the WHOLE K extent with plain 8-byte loads (16 in flight per lane, zero fill outside
// the matrix) straight into the swizzled stage layout a TMA box would have produced, one stage per K-block; after one
// __syncthreads the consumer warps run the K loop with no barrier at all.  Same tile, same K order, same DMMA sequence
// as the ring kernel: bit-identical results.  (Measured on ALIGNED operands the direct load is no faster than the ring
// -- 250^3: 8.2 us either way -- so aligned products keep the ring; profiles/r02_small_shapes.md.)
template <bool AK, bool BKM, typename TL, int REL, bool BETA, bool DIRECT = false>
__global__ void __launch_bounds__(DIRECT ? kDirectThreads : TL::THREADS, DIRECT ? 1 : TL::MIN_CTAS)
dgemm_dmma_kernel(const __grid_constant__ CUtensorMap mapA, const __grid_constant__ CUtensorMap mapB,
                  const DgemmParams p)
{
    constexpr int MI = TL::MI, NJ = TL::NJ, BM = TL::BM, BN = TL::BN, STAGES = TL::STAGES,
                  A_BYTES = TL::A_BYTES, STAGE_BYTES = TL::STAGE_BYTES, CONSUMER_WARPS = TL::CONSUMER_WARPS;
    extern __shared__ uint8_t smem_raw[];
    const uint32_t smem_base = (tma::smem_u32(smem_raw) + 1023u) & ~1023u;
    const uint32_t bar_base = smem_base + STAGES * STAGE_BYTES;      // full[STAGES], empty[STAGES]
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

    int tile_m, tile_n, tile = blockIdx.x, pa = -1, pb = -1;        // pa / pb: batch plane of A / B (-1: 2-D maps)
    int kb_begin = 0, kblocks = (int)((p.k + BK - 1) / BK);
    if (p.tiles_per_batch > 0) {
        const int b = tile / p.tiles_per_batch;
        tile -= b * p.tiles_per_batch;
        if (p.kb_per_slice > 0) {            // split-K: b is the K slice, the operands are the same 2-D arrays
            kb_begin = b * p.kb_per_slice;
            kblocks = max(0, min(p.kb_per_slice, kblocks - kb_begin));
        } else {
            pa = b * p.plane_a; pb = b * p.plane_b;
        }
    }
    tile_coords(p, tile, tile_m, tile_n);
    const int m0 = tile_m * BM, n0 = tile_n * BN;

    if constexpr (DIRECT) {
        constexpr int NW = kDirectThreads / 32, U = 16;
        const int kpad = kblocks * BK;
        // (no "memory" clobber: the next batch's loads may be scheduled above these stores; the __syncthreads below orders
        //  them against the consumers' ld.shared)
        auto sts = [](uint32_t addr, double v) { asm volatile("st.shared.f64 [%0], %1;" ::"r"(addr), "d"(v)); };
        // K-major panel (rows = the m | n index, k contiguous): (r, k) -> stage k / 16, row r, chunk ((k % 16) / 2) ^ (r % 8)
        auto load_kmajor = [&](const double *__restrict__ G, int64_t ld, int r0, int64_t rmax, int rows, uint32_t off) {
            for (int r = warp; r < rows; r += NW) {
                const bool rin = r0 + r < rmax;
                const double *row = G + (int64_t)(r0 + r) * ld;
                for (int kq = lane; kq < kpad; kq += 32 * U) {
                    double v[U];
#pragma unroll
                    for (int u = 0; u < U; ++u) {
                        const int k = kq + 32 * u;
                        v[u] = (rin && k < p.k) ? __ldg(row + k) : 0.0;
                    }
#pragma unroll
                    for (int u = 0; u < U; ++u) {
                        const int k = kq + 32 * u, kk = k & 15;
                        if (k < kpad)
                            sts(smem_base + (k >> 4) * STAGE_BYTES + off + r * 128 + ((((kk >> 1) ^ r) & 7) << 4) + (kk & 1) * 8, v[u]);
                    }
                }
            }
        };
        // MN-major panel (rows = k, the m | n index contiguous): (k, c) -> stage k / 16, box c / 16, row k % 16,
        // chunk ((c % 16) / 2) ^ (k % 8)
        auto load_mnmajor = [&](const double *__restrict__ G, int64_t ld, int c0, int64_t cmax, int cols, uint32_t off) {
            for (int c = lane; c < cols; c += 32) {
                const bool cin = c0 + c < cmax;
                const double *col = G + c0 + c;
                for (int kq = warp; kq < kpad; kq += NW * U) {
                    double v[U];
#pragma unroll
                    for (int u = 0; u < U; ++u) {
                        const int k = kq + NW * u;
                        v[u] = (cin && k < p.k) ? __ldg(col + (int64_t)k * ld) : 0.0;
                    }
#pragma unroll
                    for (int u = 0; u < U; ++u) {
                        const int k = kq + NW * u, kk = k & 15;
                        if (k < kpad)
                            sts(smem_base + (k >> 4) * STAGE_BYTES + off + (c >> 4) * 2048 + kk * 128 +
                                    (((((c & 15) >> 1) ^ kk) & 7) << 4) + (c & 1) * 8, v[u]);
                    }
                }
            }
        };
        if (AK) load_kmajor(p.Ag, p.lda, m0, p.m, BM, 0); else load_mnmajor(p.Ag, p.lda, m0, p.m, BM, 0);
        if (BKM) load_kmajor(p.Bg, p.ldb, n0, p.n, BN, A_BYTES); else load_mnmajor(p.Bg, p.ldb, n0, p.n, BN, A_BYTES);
        __syncthreads();
        if (warp >= CONSUMER_WARPS) return;
    } else {
        if (threadIdx.x == 0) {
            for (int s = 0; s < STAGES; ++s) {
                tma::mbar_init(bar_base + 8 * s, 1);                       // full: producer's expect_tx arrive
                tma::mbar_init(bar_base + 8 * (STAGES + s), CONSUMER_WARPS); // empty: one arrive per consumer warp
            }
            tma::fence_barrier_init();
        }
        __syncthreads();

        if (warp == CONSUMER_WARPS) {
            // ================================ producer ================================
            if (lane == 0) {
                tma::prefetch_tmap(&mapA);
                tma::prefetch_tmap(&mapB);
                // (L2 eviction hints -- evict_last on A, evict_first on B, tma::load_2d_hint -- were measured:
                //  DRAM traffic 6.9 GB per 8192^3 launch with or without them, time unchanged; left out.)
                // Successive waves of CTAs (one CTA per SM: blockIdx / gridDim-independent SM count) walk K in
                // opposite directions: a wave ends with the K-tail of its panels in L2, and the next wave -- which
                // shares its A panels under the grouped rasterisation -- starts on exactly that tail.
                const bool rev = p.snake_waves > 0 && ((blockIdx.x / p.snake_waves) & 1);
                for (int kb = 0; kb < kblocks; ++kb) {
                    const int s = kb % STAGES;
                    const uint32_t round = kb / STAGES;
                    if (kb >= STAGES) tma::mbar_wait(bar_base + 8 * (STAGES + s), (round - 1) & 1);
                    const uint32_t full = bar_base + 8 * s;
                    const uint32_t sa = smem_base + s * STAGE_BYTES, sb = sa + A_BYTES;
                    tma::mbar_expect_tx(full, STAGE_BYTES);
                    const int k0 = (kb_begin + (rev ? kblocks - 1 - kb : kb)) * BK;
                    if (AK) tma::load_box(sa, &mapA, full, k0, m0, pa);
                    else if (p.a_one_box) tma::load_3d(sa, &mapA, full, 0, k0, m0 >> 4);
                    else
    #pragma unroll
                        for (int j = 0; j < BM / 16; ++j) tma::load_box(sa + j * 2048, &mapA, full, m0 + 16 * j, k0, pa);
                    if (BKM) tma::load_box(sb, &mapB, full, k0, n0, pb);
                    else if (p.b_one_box) tma::load_3d(sb, &mapB, full, 0, k0, n0 >> 4);
                    else
    #pragma unroll
                        for (int j = 0; j < BN / 16; ++j) tma::load_box(sb + j * 2048, &mapB, full, n0 + 16 * j, k0, pb);
                }
            }
            return;
        }
    }

    // ================================ consumers ================================
    const int g = lane >> 2, t = lane & 3;
    const int wm = warp / TL::WN, wn = warp % TL::WN;  // WM x WN warps, warp tile (8 MI) x (8 NJ)
    // per-lane byte offsets of the four k-steps inside a stage tile (frag / box offsets are added
    // as immediates below)
    uint32_t offA[4], offB[4];
#pragma unroll
    for (int s = 0; s < 4; ++s) {
        const int kk = kslot(s, t);
        const uint32_t kmaj = g * 128 + ((((kk >> 1) ^ g) & 7) << 4) + (kk & 1) * 8;
        const uint32_t mnmaj = kk * 128 + ((((g >> 1) ^ (kk & 7)) & 7) << 4) + (g & 1) * 8;
        offA[s] = AK ? (uint32_t)(wm * MI * 8 * 128) + kmaj : (uint32_t)(wm * (MI / 2) * 2048) + mnmaj;
        offB[s] = A_BYTES + (BKM ? (uint32_t)(wn * NJ * 8 * 128) + kmaj : (uint32_t)(wn * (NJ / 2) * 2048) + mnmaj);
    }

    double acc[MI][NJ][2];
#pragma unroll
    for (int i = 0; i < MI; ++i)
#pragma unroll
        for (int j = 0; j < NJ; ++j) acc[i][j][0] = acc[i][j][1] = 0.0;

    auto load_frags = [&](uint32_t stage, int s, double (&a)[MI], double (&b)[NJ]) {
#pragma unroll
        for (int i = 0; i < MI; ++i)
            a[i] = lds64(stage + (AK ? offA[s] + i * 1024 : (offA[s] ^ ((i & 1) << 6)) + (i >> 1) * 2048));
#pragma unroll
        for (int j = 0; j < NJ; ++j)
            b[j] = lds64(stage + (BKM ? offB[s] + j * 1024 : (offB[s] ^ ((j & 1) << 6)) + (j >> 1) * 2048));
    };

    for (int kb = 0; kb < kblocks; ++kb) {
        const int s = DIRECT ? kb : kb % STAGES;
        const uint32_t stage = smem_base + s * STAGE_BYTES;
        if constexpr (!DIRECT) {
            tma::mbar_wait(bar_base + 8 * s, (kb / STAGES) & 1);
            release_before<REL, STAGES>(bar_base + 8 * STAGES, s, kb == 0, lane);
        }
        double a[2][MI], b[2][NJ];
        load_frags(stage, 0, a[0], b[0]);
#pragma unroll
        for (int st = 0; st < 4; ++st) {
            if (st < 3) load_frags(stage, st + 1, a[(st + 1) & 1], b[(st + 1) & 1]);
#pragma unroll
            for (int i = 0; i < MI; ++i)
#pragma unroll
                for (int j = 0; j < NJ; ++j) dmma(acc[i][j][0], acc[i][j][1], a[st & 1][i], b[st & 1][j]);
        }
        if constexpr (!DIRECT) release_after<REL>(bar_base + 8 * (STAGES + s), lane);   // stage consumed by this warp
    }

    // ---- epilogue: C = alpha * acc (+ beta * Cin); C itself is never read unless it is Cin ----
    // (the batch's C offset is recomputed here rather than carried through the K loop)
    dgemm_epilogue<TL, BETA>(p, acc, m0, n0, wm, wn, g, t,
                             p.tiles_per_batch > 0 ? (int64_t)(blockIdx.x / p.tiles_per_batch) * p.stride_c : 0);
}


// =====================================================================================
// Persistent variant of the 128 x 128 real kernel: ONE CTA per SM works through
//   (1) `dp_rounds` whole tiles, tile = round * gridDim + cta  (the same waves the data-parallel launch forms; the
//       producer warp runs ahead into the next tile while the consumers store the previous one), then
//   (2) its share of the STREAM-K region: the (tile, K-block) iteration space of the remaining tiles
//       [sk_tile0, tiles) cut into gridDim equal contiguous ranges -- so a product whose tile count is not a
//       multiple of the SM count (dgemm 8192^3: 4096 tiles = 27.68 waves of 148; 2048^3: 256 tiles) still loads
//       every SM evenly instead of running a last wave that is 68 % (or 73 %) full.
// A stream-K range is a sequence of segments (one tile, K-blocks [ks, ke)):
//   whole tile  (ks == 0, ke == kblocks)  -> the usual epilogue;
//   tile tail   (ks > 0): only ever the FIRST segment of a CTA -> accumulators go to the CTA's workspace
//                slot in fragment layout, then a release flag;
//   tile head   (ks == 0, ke < kblocks): only ever the LAST segment of a CTA -> waits for the flags of the
//                following CTAs that hold the rest of the tile (work they did at the very start of their
//                stream-K phase, so the wait is short and cannot chain), adds their partials in CTA order
//                (deterministic, bit-reproducible), runs the epilogue.
// Progress: CTA ids are TICKETS drawn from a global counter at start-up, not blockIdx, so id order is start
// order whatever the hardware's dispatch order.  A CTA only ever waits for HIGHER tickets' first stream-K segment,
// which needs nothing from anyone.  If some CTAs are not resident yet (SMs shared with another stream's kernel, MPS,
// a partitioned GPU), every resident CTA with a lower ticket than the first missing one can finish -- its
// dependencies are resident -- and frees its SM for the missing ones: the wait cannot deadlock, it can only take
// longer, so there is no watchdog trap in the spin.
// =====================================================================================
struct StreamKParams {
    DgemmParams p;
    int kblocks;
    int dp_rounds;          // whole tiles per CTA ahead of the stream-K region
    int sk_tile0;           // first tile of the stream-K region = dp_rounds * gridDim.x
    long long total;        // stream-K iterations = (tiles - sk_tile0) * kblocks
    int base, rem;          // CTA c owns base + (c < rem) of them starting at c * base + min(c, rem)
    double *ws;             // gridDim.x slots of 128 x 128 doubles
    unsigned *flags;        // gridDim.x words + the ticket counter at [gridDim.x], zeroed before the launch
};

__device__ __forceinline__ unsigned ld_acquire_gpu(const unsigned *p) {
    unsigned v;
    asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_release_gpu(unsigned *p, unsigned v) {
    asm volatile("st.release.gpu.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}

// Segment `seg` of a CTA (seg counts whole-tile rounds first, then `it` walks the CTA's stream-K range).  Pure function
// of (seg, it): the consumers evaluate it before AND after the K loop instead of keeping tile / ks / ke live across it
// (the loop body needs all 168 registers the 9-warp CTA can have).
__device__ __forceinline__ bool segment_at(const StreamKParams &q, int cta, int grid, int seg, int it, int it_end,
                                           int &tile, int &ks, int &ke) {
    if (seg < q.dp_rounds) {
        tile = seg * grid + cta; ks = 0; ke = q.kblocks;
        return true;
    }
    if (it >= it_end) return false;
    const int tl = it / q.kblocks;
    ks = it - tl * q.kblocks;
    ke = min(q.kblocks, ks + (it_end - it));
    tile = q.sk_tile0 + tl;
    return true;
}

template <bool AK, bool BKM, int REL>
__global__ void __launch_bounds__(Tile128x128::THREADS, 1)
dgemm_dmma_streamk_kernel(const __grid_constant__ CUtensorMap mapA, const __grid_constant__ CUtensorMap mapB,
                          const StreamKParams q)
{
    using TL = Tile128x128;
    constexpr int MI = TL::MI, NJ = TL::NJ, BM = TL::BM, BN = TL::BN, STAGES = TL::STAGES,
                  A_BYTES = TL::A_BYTES, STAGE_BYTES = TL::STAGE_BYTES, CONSUMER_WARPS = TL::CONSUMER_WARPS;
    const DgemmParams &p = q.p;
    extern __shared__ uint8_t smem_raw[];
    const uint32_t smem_base = (tma::smem_u32(smem_raw) + 1023u) & ~1023u;
    const uint32_t bar_base = smem_base + STAGES * STAGE_BYTES;
    const uint32_t ticket_slot = bar_base + 16 * STAGES;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int kblocks = q.kblocks;
    const int grid = gridDim.x;

    if (threadIdx.x == 0) {
        for (int s = 0; s < STAGES; ++s) {
            tma::mbar_init(bar_base + 8 * s, 1);
            tma::mbar_init(bar_base + 8 * (STAGES + s), CONSUMER_WARPS);
        }
        tma::fence_barrier_init();
        const unsigned ticket = atomicAdd(q.flags + grid, 1u);
        asm volatile("st.shared.u32 [%0], %1;" ::"r"(ticket_slot), "r"(ticket) : "memory");
    }
    __syncthreads();
    int cta;
    asm volatile("ld.shared.u32 %0, [%1];" : "=r"(cta) : "r"(ticket_slot));
    const int it_begin = cta * q.base + min(cta, q.rem);           // host guarantees total < 2^31
    const int it_end = it_begin + q.base + (cta < q.rem ? 1 : 0);

    if (warp == CONSUMER_WARPS) {
        // ================================ producer ================================
        if (lane == 0) {
            tma::prefetch_tmap(&mapA);
            tma::prefetch_tmap(&mapB);
            int g = 0;                                            // K-blocks issued so far (ring position)
            int it = it_begin, tile, ks, ke;
            for (int seg = 0; segment_at(q, cta, grid, seg, it, it_end, tile, ks, ke); ++seg) {
                if (seg >= q.dp_rounds) it += ke - ks;
                int tile_m, tile_n;
                tile_coords(p, tile, tile_m, tile_n);
                const int m0 = tile_m * BM, n0 = tile_n * BN;
                const bool rev = p.snake_waves > 0 && ((tile / p.snake_waves) & 1);
                for (int kb = ks; kb < ke; ++kb, ++g) {
                    const int s = g % STAGES;
                    if (g >= STAGES) tma::mbar_wait(bar_base + 8 * (STAGES + s), ((g / STAGES) - 1) & 1);
                    const uint32_t full = bar_base + 8 * s;
                    const uint32_t sa = smem_base + s * STAGE_BYTES, sb = sa + A_BYTES;
                    tma::mbar_expect_tx(full, STAGE_BYTES);
                    const int k0 = (rev ? kblocks - 1 - kb : kb) * BK;
                    if (AK) tma::load_2d(sa, &mapA, full, k0, m0);
                    else if (p.a_one_box) tma::load_3d(sa, &mapA, full, 0, k0, m0 >> 4);
                    else
#pragma unroll
                        for (int j = 0; j < BM / 16; ++j) tma::load_2d(sa + j * 2048, &mapA, full, m0 + 16 * j, k0);
                    if (BKM) tma::load_2d(sb, &mapB, full, k0, n0);
                    else if (p.b_one_box) tma::load_3d(sb, &mapB, full, 0, k0, n0 >> 4);
                    else
#pragma unroll
                        for (int j = 0; j < BN / 16; ++j) tma::load_2d(sb + j * 2048, &mapB, full, n0 + 16 * j, k0);
                }
            }
        }
        return;
    }
    // ================================ consumers ================================
    const int gq = lane >> 2, t = lane & 3;
    const int wm = warp / TL::WN, wn = warp % TL::WN;
    uint32_t offA[4], offB[4];
#pragma unroll
    for (int s = 0; s < 4; ++s) {
        const int kk = kslot(s, t);
        const uint32_t kmaj = gq * 128 + ((((kk >> 1) ^ gq) & 7) << 4) + (kk & 1) * 8;
        const uint32_t mnmaj = kk * 128 + ((((gq >> 1) ^ (kk & 7)) & 7) << 4) + (gq & 1) * 8;
        offA[s] = AK ? (uint32_t)(wm * MI * 8 * 128) + kmaj : (uint32_t)(wm * (MI / 2) * 2048) + mnmaj;
        offB[s] = A_BYTES + (BKM ? (uint32_t)(wn * NJ * 8 * 128) + kmaj : (uint32_t)(wn * (NJ / 2) * 2048) + mnmaj);
    }
    auto load_frags = [&](uint32_t stage, int s, double (&a)[MI], double (&b)[NJ]) {
#pragma unroll
        for (int i = 0; i < MI; ++i)
            a[i] = lds64(stage + (AK ? offA[s] + i * 1024 : (offA[s] ^ ((i & 1) << 6)) + (i >> 1) * 2048));
#pragma unroll
        for (int j = 0; j < NJ; ++j)
            b[j] = lds64(stage + (BKM ? offB[s] + j * 1024 : (offB[s] ^ ((j & 1) << 6)) + (j >> 1) * 2048));
    };
    constexpr int SLOT = BM * BN;                                  // doubles per workspace slot
    // ring position: stage index and phase parity, advanced incrementally (no division in the K loop)
    int rs = 0;
    uint32_t rround = 0;                                           // times the ring has wrapped: phase parity = rround & 1
    int it = it_begin;
    for (int seg = 0;; ++seg) {
        int nkb;
        {
            int tile, ks, ke;
            if (!segment_at(q, cta, grid, seg, it, it_end, tile, ks, ke)) break;
            nkb = ke - ks;
        }
        double acc[MI][NJ][2];
#pragma unroll
        for (int i = 0; i < MI; ++i)
#pragma unroll
            for (int j = 0; j < NJ; ++j) acc[i][j][0] = acc[i][j][1] = 0.0;
        for (; nkb > 0; --nkb) {
            const uint32_t stage = smem_base + rs * STAGE_BYTES;
            tma::mbar_wait(bar_base + 8 * rs, rround & 1u);
            release_before<REL, STAGES>(bar_base + 8 * STAGES, rs, (rs | rround) == 0, lane);
            double a[2][MI], b[2][NJ];
            load_frags(stage, 0, a[0], b[0]);
#pragma unroll
            for (int st = 0; st < 4; ++st) {
                if (st < 3) load_frags(stage, st + 1, a[(st + 1) & 1], b[(st + 1) & 1]);
#pragma unroll
                for (int i = 0; i < MI; ++i)
#pragma unroll
                    for (int j = 0; j < NJ; ++j) dmma(acc[i][j][0], acc[i][j][1], a[st & 1][i], b[st & 1][j]);
            }
            release_after<REL>(bar_base + 8 * (STAGES + rs), lane);
            if (++rs == STAGES) { rs = 0; ++rround; }
        }
        // the segment again (see segment_at), with the CTA's ticket re-read from shared memory
        int cta2;
        asm volatile("ld.shared.u32 %0, [%1];" : "=r"(cta2) : "r"(ticket_slot));
        const int end2 = cta2 * q.base + min(cta2, q.rem) + q.base + (cta2 < q.rem ? 1 : 0);
        int tile, ks, ke;
        segment_at(q, cta2, grid, seg, it, end2, tile, ks, ke);
        if (seg >= q.dp_rounds) it += ke - ks;
        if (ks > 0) {
            // ---- tile tail: hand the partial sums to the CTA that owns the head of this tile ----
            double2 *slot = reinterpret_cast<double2 *>(q.ws + (size_t)cta2 * SLOT);
#pragma unroll
            for (int i = 0; i < MI; ++i)
#pragma unroll
                for (int j = 0; j < NJ; ++j)
                    slot[((warp * MI + i) * NJ + j) * 32 + lane] = make_double2(acc[i][j][0], acc[i][j][1]);
            __threadfence();
            asm volatile("bar.sync 1, %0;" ::"n"(CONSUMER_WARPS * 32) : "memory");
            if (threadIdx.x == 0) st_release_gpu(q.flags + cta2, 1u);
        } else {
            if (ke < kblocks) {
                // ---- tile head: add the partial sums of the CTAs holding K-blocks [ke, kblocks) ----
                int pos = ke;
                for (int c2 = cta2 + 1; pos < kblocks; ++c2) {
                    unsigned ns = 32;
                    while (ld_acquire_gpu(q.flags + c2) == 0u) {      // no watchdog: see the progress argument above
                        __nanosleep(ns);
                        if (ns < 2048) ns *= 2;
                    }
                    const double2 *slot = reinterpret_cast<const double2 *>(q.ws + (size_t)c2 * SLOT);
#pragma unroll
                    for (int i = 0; i < MI; ++i)
#pragma unroll
                        for (int j = 0; j < NJ; ++j) {
                            const double2 v = __ldcg(&slot[((warp * MI + i) * NJ + j) * 32 + lane]);
                            acc[i][j][0] += v.x;
                            acc[i][j][1] += v.y;
                        }
                    pos += min(q.base + (c2 < q.rem ? 1 : 0), kblocks - pos);
                }
            }
            int tile_m, tile_n;
            tile_coords(p, tile, tile_m, tile_n);
            dgemm_epilogue<TL, true, 0>(p, acc, tile_m * BM, tile_n * BN, wm, wn, gq, t);
        }
    }
}

// =====================================================================================
// complex f64 (zgemm): "4M in registers".  Interleaved (re,im) operands are used as stored:
// one ld.shared.v2.f64 brings (re, im) of a fragment element, and each complex tile product is
// four DMMAs   Cr += Ar Br;  Cr += (-Ai) Bi;  Ci += Ar Bi;  Ci += Ai Br
// with the conjugations of `H` folded into the sign flips (integer XOR on the high word).
// The real and imaginary accumulators of an element live in the same lane, so the epilogue
// writes interleaved complex numbers with 16-byte stores and no planar pass exists anywhere.
//   CTA tile 128 x 64 (complex), K-block 8 complex (128 bytes), 8-stage ring of 24 KiB stages,
//   warps 4 (m) x 2 (n), warp tile 32 x 32 = 4 x 4 complex DMMA tiles (64 accumulator doubles).
//   k slots: step 0 reads k = 0,2,4,6, step 1 reads k = 1,3,5,7 (conflict-free quarter-warp
//   ld.shared.v2 under the 128-byte swizzle for both majors).
// =====================================================================================
constexpr int ZBK = 8;
// Complex tiles (extents in complex elements), WM x WN consumer warps of MI x NJ complex DMMA tiles:
//   128 x 64 (4x2 warps of 32 x 32): throughput shape; 2 consumer warpgroups + 1 producer warpgroup with
//            setmaxnreg (which works per warpgroup), 8 stages;
//    64 x 64 (2x2 warps of 32 x 32) and 32 x 32 (2x2 warps of 16 x 16): small products spread over more SMs
//            (see the real tiles above); one producer warp, no register re-partitioning needed.
template <int MI_, int NJ_, int WM_, int WN_, int STAGES_, int MIN_CTAS_, bool REGSPLIT_> struct ZTile {
    static constexpr int MI = MI_, NJ = NJ_, WM = WM_, WN = WN_;
    static constexpr int BM = 8 * MI * WM, BN = 8 * NJ * WN;
    static constexpr int CONSUMER_WARPS = WM * WN;
    static constexpr bool REGSPLIT = REGSPLIT_;
    static constexpr int THREADS = (CONSUMER_WARPS + (REGSPLIT ? 4 : 1)) * 32;
    static constexpr int STAGES = STAGES_;
    static constexpr int A_BYTES = BM * ZBK * 16, B_BYTES = BN * ZBK * 16, STAGE_BYTES = A_BYTES + B_BYTES;
    static constexpr int SMEM_BYTES = STAGES * STAGE_BYTES + 1024 + 2 * STAGES * 8;
    static constexpr int MIN_CTAS = MIN_CTAS_;
    static_assert(!REGSPLIT || CONSUMER_WARPS % 4 == 0, "setmaxnreg needs whole warpgroups");
};
using ZTile128x64 = ZTile<4, 4, 4, 2, 8, 1, true>;
using ZTile64x64  = ZTile<4, 4, 2, 2, 6, 1, false>;
using ZTile32x32  = ZTile<2, 2, 2, 2, 4, 5, false>;

__device__ __forceinline__ double2 lds128(uint32_t addr) {
    double2 v;
    asm volatile("ld.shared.v2.f64 {%0,%1}, [%2];" : "=d"(v.x), "=d"(v.y) : "r"(addr));
    return v;
}
__device__ __forceinline__ double flip(double x) {
    return __hiloint2double(__double2hiint(x) ^ 0x80000000, __double2loint(x));
}

struct ZgemmParams {
    int64_t m, n, k, ldc;
    Alpha<double2> alpha;
    BetaC<double2> beta;
    double2 *C;
    int tiles_m, tiles_n;
    int tiles_per_batch, plane_a, plane_b;   // strided batch, as in DgemmParams (0: one product, 2-D maps)
    int64_t stride_c;
    int kb_per_slice;                        // split-K, as in DgemmParams (0: whole K)
};

template <bool AK, bool BKM, bool CONJA, bool CONJB, typename TL, int REL, bool BETA>
__global__ void __launch_bounds__(TL::THREADS, TL::MIN_CTAS)
zgemm_dmma_kernel(const __grid_constant__ CUtensorMap mapA, const __grid_constant__ CUtensorMap mapB,
                  const ZgemmParams p)
{
    constexpr int MI = TL::MI, NJ = TL::NJ, ZBM = TL::BM, ZBN = TL::BN, ZSTAGES = TL::STAGES,
                  ZA_BYTES = TL::A_BYTES, ZSTAGE_BYTES = TL::STAGE_BYTES, CONSUMER_WARPS = TL::CONSUMER_WARPS;
    extern __shared__ uint8_t smem_raw[];
    const uint32_t smem_base = (tma::smem_u32(smem_raw) + 1023u) & ~1023u;
    const uint32_t bar_base = smem_base + ZSTAGES * ZSTAGE_BYTES;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

    constexpr int GROUP = 16;
    int tile_m, tile_n, pa = -1, pb = -1;                          // pa / pb: batch plane of A / B (-1: 2-D maps)
    int kb_begin = 0, kblocks = (int)((p.k + ZBK - 1) / ZBK);
    {
        int tile = blockIdx.x;
        if (p.tiles_per_batch > 0) {
            const int b = tile / p.tiles_per_batch;
            tile -= b * p.tiles_per_batch;
            if (p.kb_per_slice > 0) {        // split-K: b is the K slice
                kb_begin = b * p.kb_per_slice;
                kblocks = max(0, min(p.kb_per_slice, kblocks - kb_begin));
            } else {
                pa = b * p.plane_a; pb = b * p.plane_b;
            }
        }
        const int per_group = GROUP * p.tiles_n;
        const int grp = tile / per_group;
        const int first_m = grp * GROUP;
        const int gsz = min(p.tiles_m - first_m, GROUP);
        const int in_grp = tile - grp * per_group;
        tile_m = first_m + in_grp % gsz;
        tile_n = in_grp / gsz;
    }
    const int m0 = tile_m * ZBM, n0 = tile_n * ZBN;

    if (threadIdx.x == 0) {
        for (int s = 0; s < ZSTAGES; ++s) {
            tma::mbar_init(bar_base + 8 * s, 1);
            tma::mbar_init(bar_base + 8 * (ZSTAGES + s), CONSUMER_WARPS);
        }
        tma::fence_barrier_init();
    }
    __syncthreads();

    if (warp >= CONSUMER_WARPS) {
        // producer warpgroup: hand registers to the consumers (168 at launch -> 40 here, 232 there)
        if constexpr (TL::REGSPLIT) asm volatile("setmaxnreg.dec.sync.aligned.u32 40;");
        if (warp == CONSUMER_WARPS && lane == 0) {
            tma::prefetch_tmap(&mapA);
            tma::prefetch_tmap(&mapB);
            for (int kb = 0; kb < kblocks; ++kb) {
                const int s = kb % ZSTAGES;
                const uint32_t round = kb / ZSTAGES;
                if (kb >= ZSTAGES) tma::mbar_wait(bar_base + 8 * (ZSTAGES + s), (round - 1) & 1);
                const uint32_t full = bar_base + 8 * s;
                const uint32_t sa = smem_base + s * ZSTAGE_BYTES, sb = sa + ZA_BYTES;
                tma::mbar_expect_tx(full, ZSTAGE_BYTES);
                const int k0 = (kb_begin + kb) * ZBK;
                // tensor maps are over doubles: inner coordinates are 2 x the complex index
                if (AK) tma::load_box(sa, &mapA, full, 2 * k0, m0, pa);
                else
#pragma unroll
                    for (int j = 0; j < ZBM / 8; ++j) tma::load_box(sa + j * 1024, &mapA, full, 2 * (m0 + 8 * j), k0, pa);
                if (BKM) tma::load_box(sb, &mapB, full, 2 * k0, n0, pb);
                else
#pragma unroll
                    for (int j = 0; j < ZBN / 8; ++j) tma::load_box(sb + j * 1024, &mapB, full, 2 * (n0 + 8 * j), k0, pb);
            }
        }
        return;
    }

    if constexpr (TL::REGSPLIT) asm volatile("setmaxnreg.inc.sync.aligned.u32 232;");
    const int g = lane >> 2, t = lane & 3;
    const int wm = warp / TL::WN, wn = warp % TL::WN;  // WM x WN warps, warp tile (8 MI) x (8 NJ) complex
    uint32_t offA[2], offB[2];
#pragma unroll
    for (int s = 0; s < 2; ++s) {
        const int kk = 2 * t + s;
        const uint32_t kmaj = g * 128 + (((kk ^ g) & 7) << 4);
        const uint32_t mnmaj = kk * 128 + (((g ^ kk) & 7) << 4);
        offA[s] = (uint32_t)(wm * MI * 1024) + (AK ? kmaj : mnmaj);
        offB[s] = ZA_BYTES + (uint32_t)(wn * NJ * 1024) + (BKM ? kmaj : mnmaj);
    }

    double cr[MI][NJ][2], ci[MI][NJ][2];
#pragma unroll
    for (int i = 0; i < MI; ++i)
#pragma unroll
        for (int j = 0; j < NJ; ++j) cr[i][j][0] = cr[i][j][1] = ci[i][j][0] = ci[i][j][1] = 0.0;

    for (int kb = 0; kb < kblocks; ++kb) {
        const int s = kb % ZSTAGES;
        const uint32_t stage = smem_base + s * ZSTAGE_BYTES;
        tma::mbar_wait(bar_base + 8 * s, (kb / ZSTAGES) & 1);
        release_before<REL, ZSTAGES>(bar_base + 8 * ZSTAGES, s, kb == 0, lane);
        double2 a[2][MI], b[2][NJ];
#pragma unroll
        for (int i = 0; i < MI; ++i) a[0][i] = lds128(stage + offA[0] + i * 1024);
#pragma unroll
        for (int j = 0; j < NJ; ++j) b[0][j] = lds128(stage + offB[0] + j * 1024);
#pragma unroll
        for (int st = 0; st < 2; ++st) {
            if (st == 0) {
#pragma unroll
                for (int i = 0; i < MI; ++i) a[1][i] = lds128(stage + offA[1] + i * 1024);
#pragma unroll
                for (int j = 0; j < NJ; ++j) b[1][j] = lds128(stage + offB[1] + j * 1024);
            }
            double ai[MI], nai[MI], bi[NJ];
#pragma unroll
            for (int i = 0; i < MI; ++i) {       // ai = Im(op A), nai = -Im(op A): one sign flip either way
                ai[i] = CONJA ? flip(a[st][i].y) : a[st][i].y;
                nai[i] = CONJA ? a[st][i].y : flip(a[st][i].y);
            }
#pragma unroll
            for (int j = 0; j < NJ; ++j) bi[j] = CONJB ? flip(b[st][j].y) : b[st][j].y;
            // four passes over the 4 x 4 tile grid: 16 independent DMMAs between dependent ones
#pragma unroll
            for (int i = 0; i < MI; ++i)
#pragma unroll
                for (int j = 0; j < NJ; ++j) dmma(cr[i][j][0], cr[i][j][1], a[st][i].x, b[st][j].x);
#pragma unroll
            for (int i = 0; i < MI; ++i)
#pragma unroll
                for (int j = 0; j < NJ; ++j) dmma(ci[i][j][0], ci[i][j][1], a[st][i].x, bi[j]);
#pragma unroll
            for (int i = 0; i < MI; ++i)
#pragma unroll
                for (int j = 0; j < NJ; ++j) dmma(cr[i][j][0], cr[i][j][1], nai[i], bi[j]);
#pragma unroll
            for (int i = 0; i < MI; ++i)
#pragma unroll
                for (int j = 0; j < NJ; ++j) dmma(ci[i][j][0], ci[i][j][1], ai[i], b[st][j].x);
        }
        release_after<REL>(bar_base + 8 * (ZSTAGES + s), lane);
    }

    const double2 al = p.alpha.get();
    const bool acc_in = BETA && p.beta.on();              // BETA is a template parameter: see dgemm_epilogue
    double2 *const Cb = p.C + (p.tiles_per_batch > 0 ? (int64_t)(blockIdx.x / p.tiles_per_batch) * p.stride_c : 0);
#pragma unroll
    for (int i = 0; i < MI; ++i) {
        const int64_t r = m0 + wm * MI * 8 + i * 8 + g;
        if (r >= p.m) continue;
        double2 *crow = Cb + r * p.ldc;
#pragma unroll
        for (int j = 0; j < NJ; ++j) {
            const int64_t c = n0 + wn * NJ * 8 + j * 8 + 2 * t;
#pragma unroll
            for (int e = 0; e < 2; ++e) {
                if (c + e < p.n) {
                    const double xr = cr[i][j][e], xi = ci[i][j][e];
                    double2 v = make_double2(al.x * xr - al.y * xi, al.x * xi + al.y * xr);
                    if (acc_in) v = p.beta.add(v, r, c + e);
                    crow[c + e] = v;
                }
            }
        }
    }
}

// ---------------------------------------------------------------------------------
// host side (real f64)
// ---------------------------------------------------------------------------------
bool aligned16(const void *p) { return (reinterpret_cast<uintptr_t>(p) & 15u) == 0; }

// Tile choice.  Every resident CTA of an SM shares its one FP64 tensor pipe (64 FMA/clk), so a launch
// lasts as long as the busiest SM: ceil(tiles / SMs) tiles of BM x BN x k FMAs, stretched by the
// efficiency the tile shape reaches on the pipe (smaller tiles spend more issue slots on shared-memory
// fragment loads per DMMA) plus a per-tile prologue/epilogue expressed in k-equivalents.
// The efficiencies are measured (profiles/r01_dmma_tiles.jsonl).
struct TileModel { int bm, bn; double eff; double fixed_k; double eff_alone; };
// eff: pipe efficiency with the SM full of such CTAs, relative to 128 x 128 (4096^3: 3858 / 3977 / 3991 /
// 3926 us); eff_alone: with ONE such CTA on the SM (a 4-warp CTA cannot keep the pipe busy by itself).
constexpr TileModel kTileModel[4] = {
    {128, 128, 1.000, 32.0, 1.00},
    {64, 128, 0.970, 32.0, 0.97},
    {64, 64, 0.967, 32.0, 0.90},
    {32, 32, 0.975, 32.0, 0.50},
};

int pick_tile(int64_t m, int64_t n, int64_t k, double *cost_out = nullptr, int64_t batch = 1)
{
    if (g_tuning[ACCBLAS_TUNE_DMMA_TILE] >= 0 && g_tuning[ACCBLAS_TUNE_DMMA_TILE] < 4) return g_tuning[ACCBLAS_TUNE_DMMA_TILE];
    const int sms = sm_count();
    int best = 0;
    double best_cost = 0.0;
    for (int c = 0; c < 4; ++c) {
        const TileModel &t = kTileModel[c];
        const int64_t tiles = ((m + t.bm - 1) / t.bm) * ((n + t.bn - 1) / t.bn) * batch;     // a batch shares the waves
        const double per_sm = (double)((tiles + sms - 1) / sms);
        const double cost = per_sm * (double)t.bm * (double)t.bn * ((double)k + t.fixed_k) / (per_sm > 1.0 ? t.eff : t.eff_alone);
        if (c == 0 || cost < best_cost * 0.999) { best = c; best_cost = cost; }
    }
    if (cost_out) *cost_out = best_cost;
    return best;
}

// Stage-release mode (see release_stage).  Measured at 8192^3 (profiles/r02_dgemm_release_schedule.jsonl, ms, NN / NT / TN / TT):
//   data-parallel 128 x 128:   fenced 30.74 / 30.61 / 31.03 / 30.51     delayed 30.65 / 30.61 / 30.72 / 30.54
//   persistent (rounds + SK):  fenced 30.44 / 30.45 / 30.72 / 30.50     delayed 30.58 / 30.43 / 31.00 / 30.38
// (the delayed persistent variants of NN / NT spill: the loop body has no register left for the previous slot).
// Automatic choice: delayed on the data-parallel 128 x 128 tile, fenced in the persistent kernel, on the small
// multi-CTA tiles (whose 3-4 stage ring cannot spare a stage) and in the complex kernels.
// ACCBLAS_TUNE_RELEASE = 0 / 2 forces fenced / delayed everywhere.  The light release (1) is not built: it races.
int release_mode(bool dp_big_tile)
{
    const int knob = g_tuning[ACCBLAS_TUNE_RELEASE];
    if (knob == 0) return REL_FENCED;
    if (knob == 2) return REL_DELAYED;
    return dp_big_tile ? REL_DELAYED : REL_FENCED;
}

// tensor maps of one operand: K-major -> one 2-D box of BK x rows; MN-major -> one 3-D box when the extent allows
// (tma::make_mn3d), else BM/16 (BN/16) 2-D boxes of 16 x 16
bool make_operand_map(CUtensorMap *map, bool kmajor, const void *base, int64_t rows, int64_t cols, int64_t ld, int tile_extent,
                      int *one_box, int64_t planes = 0, int64_t plane_stride = 0)
{
    *one_box = 0;
    if (planes > 0)                              // strided batch: 3-D maps, boxes as in the 2-D case (no one-box panels)
        return tma::make_2d_batched(map, CU_TENSOR_MAP_DATA_TYPE_FLOAT64, 8, base, (uint64_t)rows, (uint64_t)cols,
                                    kmajor ? tile_extent : BK, 16, (uint64_t)ld * 8, (uint64_t)planes, (uint64_t)plane_stride * 8);
    if (kmajor)
        return tma::make_2d(map, CU_TENSOR_MAP_DATA_TYPE_FLOAT64, 8, base, (uint64_t)rows, (uint64_t)cols, tile_extent, 16, (uint64_t)ld * 8);
    if (g_tuning[ACCBLAS_TUNE_MN_ONEBOX] != 0 && cols % 16 == 0 &&
        tma::make_mn3d(map, CU_TENSOR_MAP_DATA_TYPE_FLOAT64, 8, base, (uint64_t)rows, (uint64_t)cols, 16, BK, tile_extent / 16, (uint64_t)ld * 8)) {
        *one_box = 1;
        return true;
    }
    return tma::make_2d(map, CU_TENSOR_MAP_DATA_TYPE_FLOAT64, 8, base, (uint64_t)rows, (uint64_t)cols, BK, 16, (uint64_t)ld * 8);
}

void fill_params(DgemmParams &p, const GemmArgs &g, int BM, int BN)
{
    p.m = g.m; p.n = g.n; p.k = g.k; p.ldc = g.ldc;
    p.alpha = make_alpha<double>(g.alpha, g.alpha_dev);
    p.beta = make_beta<double>(g);
    p.C = static_cast<double *>(g.C);
    p.tiles_m = (int)((g.m + BM - 1) / BM);
    p.tiles_n = (int)((g.n + BN - 1) / BN);
    p.group = g_tuning[ACCBLAS_TUNE_RASTER_GROUP] > 0 ? g_tuning[ACCBLAS_TUNE_RASTER_GROUP] : 16;
    p.tiles_per_batch = 0; p.plane_a = p.plane_b = 0; p.stride_c = 0;
    p.kb_per_slice = 0;
    p.Ag = p.Bg = nullptr; p.lda = p.ldb = 0;
}

template <bool AK, bool BKM, typename TL, int REL, bool BETA>
int launch_dgemm_tile(const GemmArgs &g)
{
    constexpr int BM = TL::BM, BN = TL::BN, SMEM_BYTES = TL::SMEM_BYTES;
    CUtensorMap mapA, mapB;
    DgemmParams p;
    const bool batched = g.batch > 1;
    // a batch that re-uses one operand (stride 0) maps it as a single plane and always loads plane 0
    if (!make_operand_map(&mapA, AK, g.A, g.rowsA, g.colsA, g.lda, BM, &p.a_one_box, batched ? (g.strideA ? g.batch : 1) : 0,
                          g.strideA ? g.strideA : g.rowsA * g.lda) ||
        !make_operand_map(&mapB, BKM, g.B, g.rowsB, g.colsB, g.ldb, BN, &p.b_one_box, batched ? (g.strideB ? g.batch : 1) : 0,
                          g.strideB ? g.strideB : g.rowsB * g.ldb))
        return cuda_fail(cudaErrorInvalidValue, "cuTensorMapEncodeTiled");
    static thread_local AttrOnce attr_once;
    if (attr_once.need()) {
        ACC_CUDA(cudaFuncSetAttribute(dgemm_dmma_kernel<AK, BKM, TL, REL, BETA>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_BYTES));
    }
    fill_params(p, g, BM, BN);
    p.snake_waves = g_tuning[ACCBLAS_TUNE_SNAKE_K] == 0 ? 0 : sm_count() * TL::MIN_CTAS;
    int64_t tiles = (int64_t)p.tiles_m * p.tiles_n;
    if (batched) {
        p.tiles_per_batch = (int)tiles;
        p.plane_a = g.strideA ? 1 : 0; p.plane_b = g.strideB ? 1 : 0;
        p.stride_c = g.strideC;
        p.snake_waves = 0;
        tiles *= g.batch;
    } else if (g.ksplit > 1) {
        const int kblocks = (int)((g.k + BK - 1) / BK);
        p.tiles_per_batch = (int)tiles;
        p.kb_per_slice = (kblocks + g.ksplit - 1) / g.ksplit;
        p.C = static_cast<double *>(g.ksplit_ws); p.ldc = g.n; p.stride_c = g.m * g.n;
        p.snake_waves = 0;
        tiles *= g.ksplit;
    }
    if (tiles > 0x7fffffff) return ACCBLAS_ERR_SHAPE;
    dgemm_dmma_kernel<AK, BKM, TL, REL, BETA><<<(unsigned)tiles, TL::THREADS, SMEM_BYTES, g.stream>>>(mapA, mapB, p);
    return check_launch("dgemm_dmma_kernel");
}

// The persistent kernel's schedule for a product: how many whole-tile rounds run ahead of the stream-K region.
struct StreamKPlan { bool use; int grid; int dp_rounds; };

// The persistent DP + stream-K kernel pays when a product of at least one full wave of 128 x 128 tiles leaves a
// partial last wave.  Measured round 1 as pure stream-K (profiles/r01_streamk.jsonl): 2048^3 (256 tiles, wave
// efficiency 0.865) 522 -> 510 us, 2560^3 (0.90) 1013 -> 975, 3072^3 (0.973) 1680 -> 1663; below one wave the small
// data-parallel tiles win (1024^3: 75.8 against 86.0 us) because the scratch, the flag memset and the fix-ups cost
// ~30 us.  Since round 2 only the LAST partial wave is stream-K'd (every CTA first runs floor(tiles / SMs) whole
// tiles), which limits the fix-up traffic to at most one 128 KiB partial per SM whatever the product size and makes
// the schedule worth taking at 0.988 too (4096^3, 8192^3: 27.68 waves of 148).
StreamKPlan plan_streamk(int64_t m, int64_t n, int64_t k)
{
    StreamKPlan pl{false, 0, 0};
    const int knob = g_tuning[ACCBLAS_TUNE_STREAMK];
    if (knob == 0 || g_tuning[ACCBLAS_TUNE_DMMA_TILE] >= 0) return pl;
    const int sms = sm_count();
    if (sms >= 1024) return pl;                                  // flag block is 4 KiB (SMs + the ticket word)
    const int64_t tiles = ((m + 127) / 128) * ((n + 127) / 128), kblocks = (k + BK - 1) / BK;
    if (tiles * kblocks < 4 * (int64_t)sms || kblocks < 4 || tiles * kblocks > 0x7fffffff) return pl;
    pl.grid = (int)std::min<int64_t>(sms, tiles * kblocks);
    const int64_t rem_tiles = tiles % pl.grid;
    // whole-tile rounds first; knob 2 = pure stream-K (round 1's schedule, kept for the side-by-side)
    pl.dp_rounds = knob == 2 ? 0 : (int)(tiles / pl.grid);
    if (rem_tiles == 0 && knob != 2) return pl;                  // full waves only: the data-parallel launch is the same thing
    if (knob == 1 || knob == 2) { pl.use = true; return pl; }
    // automatic: the tail must be worth a fix-up (at least 32 K-blocks per CTA) and the product at least one wave
    pl.use = tiles >= sms && k >= 512 && rem_tiles * kblocks >= 32 * (int64_t)pl.grid;
    return pl;
}

// Persistent launch (128 x 128 tiles, one CTA per SM).  Scratch: one accumulator slot per CTA + flags + ticket.
template <bool AK, bool BKM, int REL>
int launch_dgemm_streamk(const GemmArgs &g, const StreamKPlan &pl)
{
    using TL = Tile128x128;
    CUtensorMap mapA, mapB;
    StreamKParams q;
    DgemmParams &p = q.p;
    if (!make_operand_map(&mapA, AK, g.A, g.rowsA, g.colsA, g.lda, TL::BM, &p.a_one_box) ||
        !make_operand_map(&mapB, BKM, g.B, g.rowsB, g.colsB, g.ldb, TL::BN, &p.b_one_box))
        return cuda_fail(cudaErrorInvalidValue, "cuTensorMapEncodeTiled");
    constexpr int SMEM = TL::SMEM_BYTES + 16;                      // + the ticket word
    static thread_local AttrOnce attr_once;
    if (attr_once.need()) {
        ACC_CUDA(cudaFuncSetAttribute(dgemm_dmma_streamk_kernel<AK, BKM, REL>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM));
    }
    fill_params(p, g, TL::BM, TL::BN);
    const int grid = pl.grid;
    p.snake_waves = g_tuning[ACCBLAS_TUNE_SNAKE_K] == 0 ? 0 : grid;
    q.kblocks = (int)((g.k + BK - 1) / BK);
    q.dp_rounds = pl.dp_rounds;
    q.sk_tile0 = pl.dp_rounds * grid;
    q.total = ((long long)p.tiles_m * p.tiles_n - q.sk_tile0) * q.kblocks;
    q.base = (int)(q.total / grid);
    q.rem = (int)(q.total % grid);
    const size_t ws_bytes = (size_t)grid * TL::BM * TL::BN * 8;
    char *ws = nullptr;
    ACC_CUDA(scratch_alloc(reinterpret_cast<void **>(&ws), ws_bytes + 4096, g.stream));
    q.ws = reinterpret_cast<double *>(ws);
    q.flags = reinterpret_cast<unsigned *>(ws + ws_bytes);
    cudaError_t e = cudaMemsetAsync(q.flags, 0, 4096, g.stream);
    int st = ACCBLAS_OK;
    if (e == cudaSuccess) {
        dgemm_dmma_streamk_kernel<AK, BKM, REL><<<(unsigned)grid, TL::THREADS, SMEM, g.stream>>>(mapA, mapB, q);
        st = check_launch("dgemm_dmma_streamk_kernel");
    }
    const cudaError_t e2 = cudaFreeAsync(ws, g.stream);
    if (e != cudaSuccess) return cuda_fail(e, "cudaMemsetAsync (stream-K flags)");
    if (st == ACCBLAS_OK && e2 != cudaSuccess) return cuda_fail(e2, "cudaFreeAsync");
    return st;
}

// DIRECT launch: 32 x 32 tiles, whole K resident (one 8 KiB stage per K-block), 16 loader warps; the tensor maps are unused.
template <bool AK, bool BKM, bool BETA>
int launch_dgemm_direct_beta(const GemmArgs &g)
{
    using TL = Tile32x32Deep;                         // (only BM / BN / STAGE_BYTES matter: there is no ring)
    constexpr int MAX_SMEM = kDirectMaxKBlocks * TL::STAGE_BYTES + 1024;
    static thread_local AttrOnce attr_once;
    if (attr_once.need()) {
        ACC_CUDA(cudaFuncSetAttribute(dgemm_dmma_kernel<AK, BKM, TL, REL_FENCED, BETA, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, MAX_SMEM));
    }
    DgemmParams p;
    fill_params(p, g, TL::BM, TL::BN);
    p.a_one_box = p.b_one_box = 0; p.snake_waves = 0;
    p.Ag = static_cast<const double *>(g.A); p.Bg = static_cast<const double *>(g.B);
    p.lda = g.lda; p.ldb = g.ldb;
    const int64_t tiles = (int64_t)p.tiles_m * p.tiles_n;
    const int kblocks = (int)((g.k + BK - 1) / BK);
    static const CUtensorMap unused{};
    dgemm_dmma_kernel<AK, BKM, TL, REL_FENCED, BETA, true><<<(unsigned)tiles, kDirectThreads, (size_t)kblocks * TL::STAGE_BYTES + 1024, g.stream>>>(
        unused, unused, p);
    return check_launch("dgemm_dmma_kernel<direct>");
}
template <bool AK, bool BKM>
int launch_dgemm_direct(const GemmArgs &g)
{
    return has_beta(g) ? launch_dgemm_direct_beta<AK, BKM, true>(g) : launch_dgemm_direct_beta<AK, BKM, false>(g);
}

template <bool AK, bool BKM, typename TL>
int launch_dgemm_tile_rel(const GemmArgs &g, int rel)
{
    if (has_beta(g))
        return rel == REL_DELAYED ? launch_dgemm_tile<AK, BKM, TL, REL_DELAYED, true>(g) : launch_dgemm_tile<AK, BKM, TL, REL_FENCED, true>(g);
    return rel == REL_DELAYED ? launch_dgemm_tile<AK, BKM, TL, REL_DELAYED, false>(g) : launch_dgemm_tile<AK, BKM, TL, REL_FENCED, false>(g);
}

template <bool AK, bool BKM>
int launch_dgemm(const GemmArgs &g)
{
    const bool batched = g.batch > 1;
    const int tile = pick_tile(g.m, g.n, g.k, nullptr, batched ? g.batch : 1);
    const StreamKPlan pl = (g.no_persistent || batched) ? StreamKPlan{false, 0, 0} : plan_streamk(g.m, g.n, g.k);
    if (pl.use) {
        // Products of at most two whole-tile rounds (2048^3: 256 tiles) spend most of their time in the stream-K region and
        // its short segments, where the fence of every release weighs more than the delayed variant's spills: 2048^3
        // 513 -> 503 us (cuBLAS 496); 4096^3 3852 us either way; 8192^3 30.54 fenced against 30.60 delayed.
        const int knob = g_tuning[ACCBLAS_TUNE_RELEASE];
        const bool delayed = knob == 2 || (knob < 0 && pl.dp_rounds <= 2);
        return delayed ? launch_dgemm_streamk<AK, BKM, REL_DELAYED>(g, pl) : launch_dgemm_streamk<AK, BKM, REL_FENCED>(g, pl);
    }
    const int rel = release_mode(tile == 0);
    static const int deep = [] { const char *e = getenv("ACCBLAS_DMMA_DEEP"); return e ? atoi(e) : 1; }();
    switch (tile) {
    case 1:  return launch_dgemm_tile_rel<AK, BKM, Tile64x128>(g, rel);
    case 2:  return launch_dgemm_tile_rel<AK, BKM, Tile64x64>(g, rel);
    case 3: {
        // Split-K over idle SMs: a grid of at most half a wave of 32 x 32 tiles with a long K (250 x 250 x 1024: 64 CTAs
        // walking 64 K-blocks each, 19.4 us against cuBLAS's split-K 13.1) is cut into up to 16 K slices of at least 8
        // K-blocks -- one launch, slice = the batch index of the strided-batch grid -- whose alpha-scaled partials a second
        // small kernel sums in slice order (deterministic).  The reference's cubes (k <= 1000 at >= 1000 tiles, 250^3:
        // 16 K-blocks) are untouched.
        const int64_t tiles32 = ((g.m + 31) / 32) * ((g.n + 31) / 32), kblocks = (g.k + BK - 1) / BK;
        if (!batched && !has_beta(g) && g.ksplit == 0 && g_tuning[ACCBLAS_TUNE_SPLITK] != 0 && 2 * tiles32 <= sm_count() && kblocks >= 32) {
            // two small CTAs per SM overlap their latency-bound K loops; a split must save more than the reduce launch costs
            const int S = (int)std::min<int64_t>(std::min<int64_t>(2 * sm_count() / tiles32, kblocks / 8), 16);
            if (S >= 2 && kblocks * (S - 1) >= 32 * (int64_t)S) {
                GemmArgs h = g;
                h.ksplit = S;
                ACC_CUDA(scratch_alloc(&h.ksplit_ws, sizeof(double) * (size_t)S * (size_t)g.m * (size_t)g.n, g.stream));
                int st = launch_dgemm_tile_rel<AK, BKM, Tile32x32Deep>(h, rel);
                if (st == ACCBLAS_OK) st = splitk_reduce(h);
                const cudaError_t e = cudaFreeAsync(h.ksplit_ws, g.stream);
                if (st == ACCBLAS_OK && e != cudaSuccess) return cuda_fail(e, "cudaFreeAsync");
                return st;
            }
        }
        if (deep && ((g.m + 31) / 32) * ((g.n + 31) / 32) * (batched ? g.batch : 1) <= 3 * (int64_t)sm_count())
            return launch_dgemm_tile_rel<AK, BKM, Tile32x32Deep>(g, rel);
        return launch_dgemm_tile_rel<AK, BKM, Tile32x32>(g, rel);
    }
    default: return launch_dgemm_tile_rel<AK, BKM, Tile128x128>(g, rel);
    }
}

template <bool AK, bool BKM, bool CA, bool CB, typename TL, int REL, bool BETA>
int launch_zgemm_tile(const GemmArgs &g)
{
    constexpr int ZBM = TL::BM, ZBN = TL::BN;
    CUtensorMap mapA, mapB;
    // maps are over doubles (no complex TMA type): a complex row of c columns is 2c doubles
    const bool batched = g.batch > 1;
    const CUtensorMapDataType F64 = CU_TENSOR_MAP_DATA_TYPE_FLOAT64;
    bool ok;
    if (batched)
        ok = tma::make_2d_batched(&mapA, F64, 8, g.A, (uint64_t)g.rowsA, 2 * (uint64_t)g.colsA, AK ? ZBM : ZBK, 16, (uint64_t)g.lda * 16,
                                  g.strideA ? (uint64_t)g.batch : 1, (uint64_t)(g.strideA ? g.strideA : g.rowsA * g.lda) * 16) &&
             tma::make_2d_batched(&mapB, F64, 8, g.B, (uint64_t)g.rowsB, 2 * (uint64_t)g.colsB, BKM ? ZBN : ZBK, 16, (uint64_t)g.ldb * 16,
                                  g.strideB ? (uint64_t)g.batch : 1, (uint64_t)(g.strideB ? g.strideB : g.rowsB * g.ldb) * 16);
    else
        ok = tma::make_2d(&mapA, F64, 8, g.A, (uint64_t)g.rowsA, 2 * (uint64_t)g.colsA, AK ? ZBM : ZBK, 16, (uint64_t)g.lda * 16) &&
             tma::make_2d(&mapB, F64, 8, g.B, (uint64_t)g.rowsB, 2 * (uint64_t)g.colsB, BKM ? ZBN : ZBK, 16, (uint64_t)g.ldb * 16);
    if (!ok) return cuda_fail(cudaErrorInvalidValue, "cuTensorMapEncodeTiled");
    static thread_local AttrOnce attr_once;
    if (attr_once.need()) {
        ACC_CUDA(cudaFuncSetAttribute(zgemm_dmma_kernel<AK, BKM, CA, CB, TL, REL, BETA>, cudaFuncAttributeMaxDynamicSharedMemorySize, TL::SMEM_BYTES));
    }
    ZgemmParams p;
    p.m = g.m; p.n = g.n; p.k = g.k; p.ldc = g.ldc;
    p.alpha = make_alpha<double2>(g.alpha, g.alpha_dev);
    p.beta = make_beta<double2>(g);
    p.C = static_cast<double2 *>(g.C);
    p.tiles_m = (int)((g.m + ZBM - 1) / ZBM);
    p.tiles_n = (int)((g.n + ZBN - 1) / ZBN);
    int64_t tiles = (int64_t)p.tiles_m * p.tiles_n;
    p.tiles_per_batch = 0; p.plane_a = p.plane_b = 0; p.stride_c = 0;
    p.kb_per_slice = 0;
    if (batched) {
        p.tiles_per_batch = (int)tiles;
        p.plane_a = g.strideA ? 1 : 0; p.plane_b = g.strideB ? 1 : 0;
        p.stride_c = g.strideC;
        tiles *= g.batch;
    } else if (g.ksplit > 1) {
        const int kblocks = (int)((g.k + ZBK - 1) / ZBK);
        p.tiles_per_batch = (int)tiles;
        p.kb_per_slice = (kblocks + g.ksplit - 1) / g.ksplit;
        p.C = static_cast<double2 *>(g.ksplit_ws); p.ldc = g.n; p.stride_c = g.m * g.n;
        tiles *= g.ksplit;
    }
    if (tiles > 0x7fffffff) return ACCBLAS_ERR_SHAPE;
    zgemm_dmma_kernel<AK, BKM, CA, CB, TL, REL, BETA><<<(unsigned)tiles, TL::THREADS, TL::SMEM_BYTES, g.stream>>>(mapA, mapB, p);
    return check_launch("zgemm_dmma_kernel");
}

// same cost model as pick_tile; a complex tile element costs four real FMAs
constexpr TileModel kZTileModel[3] = {
    {128, 64, 1.000, 16.0, 1.00},
    {64, 64, 0.880, 16.0, 0.88},
    {32, 32, 0.975, 16.0, 0.60},
};

int pick_ztile(int64_t m, int64_t n, int64_t k, int64_t batch = 1)
{
    if (g_tuning[ACCBLAS_TUNE_ZMMA_TILE] >= 0 && g_tuning[ACCBLAS_TUNE_ZMMA_TILE] < 3) return g_tuning[ACCBLAS_TUNE_ZMMA_TILE];
    const int sms = sm_count();
    int best = 0;
    double best_cost = 0.0;
    for (int c = 0; c < 3; ++c) {
        const TileModel &t = kZTileModel[c];
        const int64_t tiles = ((m + t.bm - 1) / t.bm) * ((n + t.bn - 1) / t.bn) * batch;
        const double per_sm = (double)((tiles + sms - 1) / sms);
        const double cost = per_sm * (double)t.bm * (double)t.bn * ((double)k + t.fixed_k) / (per_sm > 1.0 ? t.eff : t.eff_alone);
        if (c == 0 || cost < best_cost * 0.999) { best = c; best_cost = cost; }
    }
    return best;
}

template <bool AK, bool BKM, bool CA, bool CB>
int launch_zgemm(const GemmArgs &g)
{
    const int tile = pick_ztile(g.m, g.n, g.k, g.batch > 1 ? g.batch : 1);
    // complex kernels: fenced or delayed only (the light release is known to race on the 32 x 32 tile)
    const bool delayed = g_tuning[ACCBLAS_TUNE_RELEASE] == 2, beta = has_beta(g);
    switch (tile) {
    case 1:  return delayed ? launch_zgemm_tile<AK, BKM, CA, CB, ZTile64x64, REL_DELAYED, true>(g)
                     : beta ? launch_zgemm_tile<AK, BKM, CA, CB, ZTile64x64, REL_FENCED, true>(g) : launch_zgemm_tile<AK, BKM, CA, CB, ZTile64x64, REL_FENCED, false>(g);
    case 2: {
        // split-K over idle SMs, as for the real 32 x 32 tile (launch_dgemm); a complex K-block is 8 wide
        const int64_t tiles32 = ((g.m + 31) / 32) * ((g.n + 31) / 32), kblocks = (g.k + ZBK - 1) / ZBK;
        if (g.batch <= 1 && !beta && !delayed && g.ksplit == 0 && g_tuning[ACCBLAS_TUNE_SPLITK] != 0 && 2 * tiles32 <= sm_count() && kblocks >= 12) {
            // a complex K-block (8 wide, four DMMAs per tile product) costs ~0.5 us of a one-wave CTA's loop against ~0.3 us for
            // a real one: the split pays from 16 K-blocks of saving, in slices of at least 4
            const int S = (int)std::min<int64_t>(std::min<int64_t>(2 * sm_count() / tiles32, kblocks / 4), 16);
            if (S >= 2 && kblocks * (S - 1) >= 16 * (int64_t)S) {
                GemmArgs h = g;
                h.ksplit = S;
                ACC_CUDA(scratch_alloc(&h.ksplit_ws, sizeof(double2) * (size_t)S * (size_t)g.m * (size_t)g.n, g.stream));
                int st = launch_zgemm_tile<AK, BKM, CA, CB, ZTile32x32, REL_FENCED, false>(h);
                if (st == ACCBLAS_OK) st = splitk_reduce(h);
                const cudaError_t e = cudaFreeAsync(h.ksplit_ws, g.stream);
                if (st == ACCBLAS_OK && e != cudaSuccess) return cuda_fail(e, "cudaFreeAsync");
                return st;
            }
        }
        return delayed ? launch_zgemm_tile<AK, BKM, CA, CB, ZTile32x32, REL_DELAYED, true>(g)
                     : beta ? launch_zgemm_tile<AK, BKM, CA, CB, ZTile32x32, REL_FENCED, true>(g) : launch_zgemm_tile<AK, BKM, CA, CB, ZTile32x32, REL_FENCED, false>(g);
    }
    default: return delayed ? launch_zgemm_tile<AK, BKM, CA, CB, ZTile128x64, REL_DELAYED, true>(g)
                     : beta ? launch_zgemm_tile<AK, BKM, CA, CB, ZTile128x64, REL_FENCED, true>(g) : launch_zgemm_tile<AK, BKM, CA, CB, ZTile128x64, REL_FENCED, false>(g);
    }
}

template <bool AK, bool BKM>
int zgemm_conj(const GemmArgs &g)
{
    // conjugation only ever comes with a transpose (op H): CA implies !AK, CB implies BKM
    const bool ca = g.opA == ACCBLAS_OP_H, cb = g.opB == ACCBLAS_OP_H;
    if constexpr (!AK && BKM) {
        if (ca && cb) return launch_zgemm<AK, BKM, true, true>(g);
        if (ca) return launch_zgemm<AK, BKM, true, false>(g);
        if (cb) return launch_zgemm<AK, BKM, false, true>(g);
    } else if constexpr (!AK) {
        if (ca) return launch_zgemm<AK, BKM, true, false>(g);
    } else if constexpr (BKM) {
        if (cb) return launch_zgemm<AK, BKM, false, true>(g);
    }
    return launch_zgemm<AK, BKM, false, false>(g);
}

}  // namespace

bool dgemm_dmma_eligible(const GemmArgs &g)
{
    if (g.dtype != ACCBLAS_F64 && g.dtype != ACCBLAS_C64) return false;
    if (g.m < 1 || g.n < 1 || g.k < 1) return false;
    // TMA: 16-byte aligned base and row pitch
    if (!aligned16(g.A) || !aligned16(g.B) || !aligned16(g.C)) return false;
    if (g.dtype == ACCBLAS_C64) {
        if (g.rowsA > 0x3fffffff || g.colsA > 0x3fffffff || g.rowsB > 0x3fffffff || g.colsB > 0x3fffffff) return false;
        return g.m * g.n >= 32 * 32 && g.k >= 4;
    }
    if ((g.lda & 1) || (g.ldb & 1)) return false;          // row pitch must be a multiple of 16 bytes
    if (g.rowsA > 0x7fffffff || g.colsA > 0x7fffffff || g.rowsB > 0x7fffffff || g.colsB > 0x7fffffff) return false;
    // tiny products stay on the SIMT kernel: a 128 x 128 tile would be mostly padding
    if (g.m * g.n < 64 * 64 || g.k < 8) return false;
    return true;
}

size_t dgemm_dmma_workspace(const GemmArgs &g)
{
    if (g.dtype != ACCBLAS_F64) return 0;
    const StreamKPlan pl = plan_streamk(g.m, g.n, g.k);
    return pl.use ? (size_t)pl.grid * 128 * 128 * 8 + 4096 : 0;                        // accumulator slots + flags
}

// f64 products with an operand off TMA's 16-byte grid (odd column count, 8-byte aligned rows) that fit the direct kernel:
// one wave of 32 x 32 tiles at one CTA per SM, whole K in shared memory, one product
bool dgemm_direct_eligible(const GemmArgs &g)
{
    if (g_tuning[ACCBLAS_TUNE_DGEMM_DIRECT] == 0 || g.dtype != ACCBLAS_F64 || g.batch > 1) return false;
    if (g.k > (int64_t)BK * kDirectMaxKBlocks || ((g.m + 31) / 32) * ((g.n + 31) / 32) > (int64_t)sm_count()) return false;
    if (!aligned16(g.C) || g.m * g.n < 24 * 24 || g.k < 4) return false;
    return g.rowsA <= 0x7fffffff && g.colsA <= 0x7fffffff && g.rowsB <= 0x7fffffff && g.colsB <= 0x7fffffff;
}
int dgemm_direct(const GemmArgs &g)
{
    const bool ak = g.opA == ACCBLAS_OP_N, bk = g.opB != ACCBLAS_OP_N;
    if (ak && bk) return launch_dgemm_direct<true, true>(g);
    if (ak && !bk) return launch_dgemm_direct<true, false>(g);
    if (!ak && bk) return launch_dgemm_direct<false, true>(g);
    return launch_dgemm_direct<false, false>(g);
}

int dgemm_dmma(const GemmArgs &g)
{
    const bool ak = g.opA == ACCBLAS_OP_N, bk = g.opB != ACCBLAS_OP_N;
    if (g.dtype == ACCBLAS_C64) {
        if (ak && bk) return zgemm_conj<true, true>(g);
        if (ak && !bk) return zgemm_conj<true, false>(g);
        if (!ak && bk) return zgemm_conj<false, true>(g);
        return zgemm_conj<false, false>(g);
    }
    if (ak && bk) return launch_dgemm<true, true>(g);
    if (ak && !bk) return launch_dgemm<true, false>(g);
    if (!ak && bk) return launch_dgemm<false, true>(g);
    return launch_dgemm<false, false>(g);
}

}  // namespace accblas
