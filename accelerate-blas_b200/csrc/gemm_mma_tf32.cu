// gemm_mma_tf32.cu -- FP32 / complex-FP32 GEMM for SMALL and MEDIUM products: one launch, no
// pre-pass, no workspace.  Replaces BLAS.sgemm / BLAS.cgemm (LLVM/PTX/Level3.hs:91,101) where the
// tcgen05 path (gemm_tf32x3.cu: pack pass + 128 x 256 tiles) is launch- and occupancy-bound --
// in particular on the reference's own benchmark shapes, 100^3 .. 1000^3
// (bench/Bench/Accelerate.hs:92-125).
//
// Same split-precision arithmetic as the tcgen05 path (FP32-class accuracy from TF32 products):
//     a = a_hi + a_lo,  a_hi = rna_tf32(a),  a_lo = rna_tf32(a - a_hi)
//     a*b ~= a_lo*b_hi + a_hi*b_lo + a_hi*b_hi
// but the split happens in REGISTERS on the fragments, so the operands are consumed as stored:
//   * TMA (128-byte swizzle, zero fill) -> smem ring -> warp-level mma.sync.m16n8k8.tf32 (SASS HMMA),
//     producer warp + consumer warps, full/empty mbarriers -- the skeleton of gemm_dmma.cu;
//   * every op is served from the stored layout: K-major operands as one box per stage, MN-major
//     operands as 32-float-wide boxes; no transpose pass;
//   * the 8 k indices of one MMA are permuted -- slots (t, t+4) of step s read
//     k = 8s + {0,3,5,6}[t] and 8s + {1,2,4,7}[t] for BOTH operands -- which makes every warp-wide
//     ld.shared.f32 conflict-free under the swizzle for K-major (needs k mod 4 distinct over t) and
//     MN-major (needs (k mod 8) / 2 distinct over t) operands alike;
//   * the accumulation chain is cut every PROMO_KB K-blocks: the MMA accumulators are added into a
//     second register set with round-to-nearest FADDs and restart from zero (the tensor core adds
//     with truncation; see the accuracy note in gemm_tf32x3.cu);
//   * complex: the real embedding of gemm_tf32x3.cu built at FRAGMENT level -- a lane's (re, im)
//     pair of A feeds slots t / t+4 of the A fragment, its (re, im) pair of B becomes the two B
//     fragments (br, -bi) and (bi, br) whose products are Re C and Im C of the same elements, so
//     results leave as interleaved float2 stores.  2 MMAs x 3 splits per 16 x 8 x 4 complex tile
//     product: exactly the 3x of the real case, nothing wasted ("4M" arithmetic).
#include "common.cuh"
#include "refit.cuh"
#include "tma.cuh"

namespace accblas {
namespace {

__device__ __forceinline__ float lds32(uint32_t addr) {
    float v;
    asm volatile("ld.shared.f32 %0, [%1];" : "=f"(v) : "r"(addr));
    return v;
}
__device__ __forceinline__ float2 lds64f(uint32_t addr) {
    float2 v;
    asm volatile("ld.shared.v2.f32 {%0,%1}, [%2];" : "=f"(v.x), "=f"(v.y) : "r"(addr));
    return v;
}
// v -> (hi, lo) TF32 planes, as 32-bit patterns for the MMA
__device__ __forceinline__ void split_tf32(float v, uint32_t &hi, uint32_t &lo) {
    asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(hi) : "f"(v));
    const float r = v - __uint_as_float(hi);
    asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(lo) : "f"(r));
}
__device__ __forceinline__ void mma_tf32(float (&d)[4], const uint32_t (&a)[4], const uint32_t (&b)[2]) {
    asm volatile(
        "mma.sync.aligned.m16n8k8.row.col.f32.tf32.tf32.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
        : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
        : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b[0]), "r"(b[1]));
}
// d += a_lo b_hi + a_hi b_lo + a_hi b_hi  (small terms first)
__device__ __forceinline__ void mma_3x(float (&d)[4], const uint32_t (&ahi)[4], const uint32_t (&alo)[4],
                                       const uint32_t (&bhi)[2], const uint32_t (&blo)[2]) {
    mma_tf32(d, alo, bhi);
    mma_tf32(d, ahi, blo);
    mma_tf32(d, ahi, bhi);
}

constexpr int PROMO_KB = 4;     // K-blocks (of 32 real / 16 complex k) per accumulation chunk
constexpr int kDirectThreads = 512;   // DIRECT kernels: 16 loader warps (the first CONSUMER_WARPS of them go on to multiply)

// CTA tiles: WM x WN consumer warps, each MI x NJ MMA tiles of 16 x 8.
template <int MI_, int NJ_, int WM_, int WN_, int STAGES_, int MIN_CTAS_> struct STile {
    static constexpr int MI = MI_, NJ = NJ_, WM = WM_, WN = WN_;
    static constexpr int BM = 16 * MI * WM, BN = 8 * NJ * WN;
    static constexpr int CONSUMER_WARPS = WM * WN;
    static constexpr int THREADS = (CONSUMER_WARPS + 1) * 32;
    static constexpr int STAGES = STAGES_;
    static constexpr int A_BYTES = BM * 128, B_BYTES = BN * 128;          // 128 bytes of k per row
    static constexpr int STAGE_BYTES = A_BYTES + B_BYTES;
    static constexpr int SMEM_BYTES = STAGES * STAGE_BYTES + 1024 + 2 * STAGES * 8;
    static constexpr int MIN_CTAS = MIN_CTAS_;
};
using STile128 = STile<4, 2, 2, 4, 4, 1>;   // 128 x  64, 8 warps of 64 x 16
using STile64  = STile<2, 4, 2, 2, 4, 3>;   //  64 x  64, 4 warps of 32 x 32
using STile32  = STile<1, 2, 2, 2, 3, 7>;   //  32 x  32, 4 warps of 16 x 16

struct SgemmMmaParams {
    int64_t m, n, k, ldc;
    Alpha<float> alpha;        // real kernels
    Alpha<float2> calpha;      // complex kernels
    BetaC<float> beta;         // + beta * Cin (accblas_gemm_ex)
    BetaC<float2> cbeta;
    float *C;
    int tiles_m, tiles_n;
    int tiles_per_batch, plane_a, plane_b;   // strided batch as in gemm_dmma.cu's DgemmParams (0: one product, 2-D maps)
    int64_t stride_c;                        // ELEMENTS (complex elements for c32) between the outputs of the batch
    int kb_per_slice;                        // split-K as in gemm_dmma.cu: the batch index is a K slice of this many K-blocks (0: whole K)
    Refit32 refit;             // non-finite sums are recomputed in plain FP32 from the caller's operands (refit.cuh)
};

// the K range [k0, k0 + len) of a product as its own refit problem (split-K: a slice recomputes ITS inner product);
// fpe = floats per element
__device__ __forceinline__ Refit32 slice_refit(Refit32 f, int k0, int len, int fpe)
{
    if (k0 == 0 && len >= f.k) return f;
    f.A += (int64_t)fpe * (f.ta ? (int64_t)k0 * f.lda : k0);
    f.B += (int64_t)fpe * (f.tb ? k0 : (int64_t)k0 * f.ldb);
    f.k = min((int64_t)len, f.k - k0);
    return f;
}

__device__ __forceinline__ void raster(const SgemmMmaParams &p, int &tile_m, int &tile_n, int &pa, int &pb)
{
    constexpr int GROUP = 16;
    const int tiles_m = p.tiles_m, tiles_n = p.tiles_n;
    int tile = blockIdx.x;
    pa = pb = -1;                                                 // batch plane of A / B (-1: 2-D maps)
    if (p.tiles_per_batch > 0) {
        const int b = tile / p.tiles_per_batch;
        tile -= b * p.tiles_per_batch;
        pa = b * p.plane_a; pb = b * p.plane_b;
    }
    const int per_group = GROUP * tiles_n;
    const int grp = tile / per_group;
    const int first_m = grp * GROUP;
    const int gsz = min(tiles_m - first_m, GROUP);
    const int in_grp = tile - grp * per_group;
    tile_m = first_m + in_grp % gsz;
    tile_n = in_grp / gsz;
}

// =====================================================================================
// real f32.  K-block = 32 floats.  AK: op(A) K-major in memory (opA == N); BKM: op(B) K-major (opB == T).
// =====================================================================================
//
// DIRECT (round 2): operands whose base or row pitch misses TMA's 16-byte rule (the reference benchmarks 250 x 250 Floats:
// 1000-byte rows) used to be copied into pitch-aligned scratch by a separate launch first (api.cu: gemm_repitched) --
// 4 us of copy kernel plus a 3 us launch gap in front of an 8 us product.  For a product whose whole K extent fits
// in shared memory the CTA instead loads ITS OWN panels with plain 4-byte loads (every warp, coalesced along the
// stored rows, zero fill outside the matrix) straight into the swizzled stage layout TMA would have produced -- one
// stage per K-block, no ring, no mbarriers -- and the consumer code below runs unchanged.  One launch, no scratch.
template <bool AK, bool BKM, typename TL, bool DIRECT = false>
__global__ void __launch_bounds__(DIRECT ? kDirectThreads : TL::THREADS, DIRECT ? 1 : TL::MIN_CTAS)
sgemm_mma_kernel(const __grid_constant__ CUtensorMap mapA, const __grid_constant__ CUtensorMap mapB,
                 const SgemmMmaParams p)
{
    constexpr int MI = TL::MI, NJ = TL::NJ, BM = TL::BM, BN = TL::BN, STAGES = TL::STAGES, BK = 32,
                  A_BYTES = TL::A_BYTES, STAGE_BYTES = TL::STAGE_BYTES, CONSUMER_WARPS = TL::CONSUMER_WARPS;
    static_assert(MI == 1 || MI % 2 == 0, "MN-major A: 16-row MMA tiles pair up inside 32-wide boxes");
    static_assert(BM % 32 == 0 && BN % 32 == 0, "MN-major operands come in 32-wide boxes");
    extern __shared__ uint8_t smem_raw[];
    const uint32_t smem_base = (tma::smem_u32(smem_raw) + 1023u) & ~1023u;
    const uint32_t bar_base = smem_base + STAGES * STAGE_BYTES;      // full[STAGES], empty[STAGES]
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    int tile_m, tile_n, pa, pb;
    raster(p, tile_m, tile_n, pa, pb);
    const int m0 = tile_m * BM, n0 = tile_n * BN;
    int kb_begin = 0, kblocks = (int)((p.k + BK - 1) / BK);
    if (p.kb_per_slice > 0) {                 // split-K: the batch index is the K slice; operands are the same 2-D arrays
        kb_begin = (int)(blockIdx.x / p.tiles_per_batch) * p.kb_per_slice;
        kblocks = max(0, min(p.kb_per_slice, kblocks - kb_begin));
        pa = pb = -1;
    }

    if constexpr (DIRECT) {
        // ---- every warp: this CTA's panels, all K-blocks, into stage kb = smem_base + kb * STAGE_BYTES ----
        const float *__restrict__ Ag = p.refit.A, *__restrict__ Bg = p.refit.B;
        const int64_t lda = p.refit.lda, ldb = p.refit.ldb;
        const int kpad = kblocks * BK;
        constexpr int NW = kDirectThreads / 32;          // loads in flight decide this phase: 16 warps x 16 x 128 bytes
        // (no "memory" clobber: the loads of the next batch may be scheduled above these stores; the __syncthreads below
        //  orders them against the consumers' ld.shared)
        auto sts = [](uint32_t addr, float v) { asm volatile("st.shared.f32 [%0], %1;" ::"r"(addr), "f"(v)); };
        constexpr int U = 16;                             // loads in flight per lane
        // K-major panel (rows = the m | n index, k contiguous): element (r, k) -> stage k / 32, row r, 16-byte chunk
        // ((k % 32) / 4) ^ (r % 8): the 128-byte swizzle of a 32-float-wide box
        auto load_kmajor = [&](const float *__restrict__ G, int64_t ld, int r0, int64_t rmax, int rows, uint32_t off) {
            for (int r = warp; r < rows; r += NW) {
                const bool rin = r0 + r < rmax;
                const float *row = G + (int64_t)(r0 + r) * ld;
                for (int kq = lane; kq < kpad; kq += 32 * U) {
                    float v[U];
#pragma unroll
                    for (int u = 0; u < U; ++u) {
                        const int k = kq + 32 * u;
                        v[u] = (rin && k < p.k) ? __ldg(row + k) : 0.f;
                    }
#pragma unroll
                    for (int u = 0; u < U; ++u) {
                        const int k = kq + 32 * u, kk = k & 31;
                        if (k < kpad)
                            sts(smem_base + (k >> 5) * STAGE_BYTES + off + r * 128 + ((((kk >> 2) ^ r) & 7) << 4) + (kk & 3) * 4, v[u]);
                    }
                }
            }
        };
        // MN-major panel (rows = k, the m | n index contiguous): element (k, c) -> stage k / 32, box c / 32, row k % 32,
        // chunk ((c % 32) / 4) ^ (k % 8).  A lane keeps its column and walks U rows of k per batch.
        auto load_mnmajor = [&](const float *__restrict__ G, int64_t ld, int c0, int64_t cmax, int cols, uint32_t off) {
            for (int c = lane; c < cols; c += 32) {
                const bool cin = c0 + c < cmax;
                const float *col = G + c0 + c;
                for (int kq = warp; kq < kpad; kq += NW * U) {
                    float v[U];
#pragma unroll
                    for (int u = 0; u < U; ++u) {
                        const int k = kq + NW * u;
                        v[u] = (cin && k < p.k) ? __ldg(col + (int64_t)k * ld) : 0.f;
                    }
#pragma unroll
                    for (int u = 0; u < U; ++u) {
                        const int k = kq + NW * u, kk = k & 31;
                        if (k < kpad)
                            sts(smem_base + (k >> 5) * STAGE_BYTES + off + (c >> 5) * 4096 + kk * 128 +
                                    (((((c & 31) >> 2) ^ kk) & 7) << 4) + (c & 3) * 4, v[u]);
                    }
                }
            }
        };
        if (AK) load_kmajor(Ag, lda, m0, p.m, BM, 0); else load_mnmajor(Ag, lda, m0, p.m, BM, 0);
        if (BKM) load_kmajor(Bg, ldb, n0, p.n, BN, A_BYTES); else load_mnmajor(Bg, ldb, n0, p.n, BN, A_BYTES);
        __syncthreads();
        if (warp >= CONSUMER_WARPS) return;
    } else {
        if (threadIdx.x == 0) {
            for (int s = 0; s < STAGES; ++s) {
                tma::mbar_init(bar_base + 8 * s, 1);
                tma::mbar_init(bar_base + 8 * (STAGES + s), CONSUMER_WARPS);
            }
            tma::fence_barrier_init();
        }
        __syncthreads();

        if (warp == CONSUMER_WARPS) {
            // ================================ producer ================================
            if (lane == 0) {
                tma::prefetch_tmap(&mapA);
                tma::prefetch_tmap(&mapB);
                for (int kb = 0; kb < kblocks; ++kb) {
                    const int s = kb % STAGES;
                    if (kb >= STAGES) tma::mbar_wait(bar_base + 8 * (STAGES + s), ((kb / STAGES) - 1) & 1);
                    const uint32_t full = bar_base + 8 * s;
                    const uint32_t sa = smem_base + s * STAGE_BYTES, sb = sa + A_BYTES;
                    tma::mbar_expect_tx(full, STAGE_BYTES);
                    const int k0 = (kb_begin + kb) * BK;
                    if (AK) tma::load_box(sa, &mapA, full, k0, m0, pa);
                    else
    #pragma unroll
                        for (int j = 0; j < BM / 32; ++j) tma::load_box(sa + j * 4096, &mapA, full, m0 + 32 * j, k0, pa);
                    if (BKM) tma::load_box(sb, &mapB, full, k0, n0, pb);
                    else
    #pragma unroll
                        for (int j = 0; j < BN / 32; ++j) tma::load_box(sb + j * 4096, &mapB, full, n0 + 32 * j, k0, pb);
                }
            }
            return;
        }
    }

    // ================================ consumers ================================
    const int g = lane >> 2, t = lane & 3;
    const int wm = warp / TL::WN, wn = warp % TL::WN;
    // slot t of half h reads k = 8 s + perm (see header)
    const int perm[2] = {(0x6530 >> (4 * t)) & 15, (0x7421 >> (4 * t)) & 15};
    uint32_t offA[2], offB[2];
#pragma unroll
    for (int h = 0; h < 2; ++h) {
        const int pk = perm[h];
        {
            const int r = wm * MI * 16 + g;                       // first row of this lane (op(A) row index)
            offA[h] = AK ? (uint32_t)(r * 128 + ((((pk >> 2) ^ g) & 7) << 4) + (pk & 3) * 4)
                         : (uint32_t)((r >> 5) * 4096 + pk * 128 + (((((r & 31) >> 2) ^ pk) & 7) << 4) + (r & 3) * 4);
        }
        {
            const int c = wn * NJ * 8 + g;                        // first column of this lane
            offB[h] = A_BYTES + (BKM ? (uint32_t)(c * 128 + ((((pk >> 2) ^ g) & 7) << 4) + (pk & 3) * 4)
                                     : (uint32_t)((c >> 5) * 4096 + pk * 128 + (((((c & 31) >> 2) ^ pk) & 7) << 4) + (c & 3) * 4));
        }
    }

    float acc[MI][NJ][4], tot[MI][NJ][4];
#pragma unroll
    for (int i = 0; i < MI; ++i)
#pragma unroll
        for (int j = 0; j < NJ; ++j)
#pragma unroll
            for (int e = 0; e < 4; ++e) acc[i][j][e] = tot[i][j][e] = 0.f;

    for (int kb = 0; kb < kblocks; ++kb) {
        const int s = DIRECT ? kb : kb % STAGES;
        const uint32_t stage = smem_base + s * STAGE_BYTES;
        if constexpr (!DIRECT) tma::mbar_wait(bar_base + 8 * s, (kb / STAGES) & 1);
#pragma unroll
        for (int st = 0; st < 4; ++st) {
            uint32_t ahi[MI][4], alo[MI][4], bhi[NJ][2], blo[NJ][2];
#pragma unroll
            for (int i = 0; i < MI; ++i)
#pragma unroll
                for (int h = 0; h < 2; ++h) {
                    // K-major: k chunk index 2 st lands as XOR (st << 5); rows +8: +1024 B; tile i: +2048 B
                    // MN-major: k row 8 st: + st * 1024; m + 8: XOR 32; tile i: XOR ((i & 1) << 6) + (i / 2) boxes
                    const uint32_t a0 = AK ? stage + (offA[h] ^ (uint32_t)(st << 5)) + i * 2048
                                           : stage + (offA[h] ^ (uint32_t)((i & 1) << 6)) + st * 1024 + (i >> 1) * 4096;
                    const uint32_t a1 = AK ? a0 + 1024
                                           : stage + (offA[h] ^ (uint32_t)(((i & 1) << 6) | 32)) + st * 1024 + (i >> 1) * 4096;
                    split_tf32(lds32(a0), ahi[i][2 * h], alo[i][2 * h]);
                    split_tf32(lds32(a1), ahi[i][2 * h + 1], alo[i][2 * h + 1]);
                }
#pragma unroll
            for (int j = 0; j < NJ; ++j)
#pragma unroll
                for (int h = 0; h < 2; ++h) {
                    const uint32_t b0 = BKM ? stage + (offB[h] ^ (uint32_t)(st << 5)) + j * 1024
                                            : stage + (offB[h] ^ (uint32_t)((j & 3) << 5)) + st * 1024 + (j >> 2) * 4096;
                    split_tf32(lds32(b0), bhi[j][h], blo[j][h]);
                }
#pragma unroll
            for (int i = 0; i < MI; ++i)
#pragma unroll
                for (int j = 0; j < NJ; ++j) mma_3x(acc[i][j], ahi[i], alo[i], bhi[j], blo[j]);
        }
        if constexpr (!DIRECT) {
            tma::consumer_release_fence();
            if (lane == 0) tma::mbar_arrive(bar_base + 8 * (STAGES + s));   // stage consumed by this warp
        }
        if ((kb + 1) % PROMO_KB == 0) {                                 // cut the truncating chain
#pragma unroll
            for (int i = 0; i < MI; ++i)
#pragma unroll
                for (int j = 0; j < NJ; ++j)
#pragma unroll
                    for (int e = 0; e < 4; ++e) { tot[i][j][e] += acc[i][j][e]; acc[i][j][e] = 0.f; }
        }
    }

    const int64_t batch = p.tiles_per_batch > 0 ? blockIdx.x / p.tiles_per_batch : 0;   // recomputed: not carried through the K loop
    // ---- non-finite sums (Inf / NaN / overflowing inputs cannot go through the split): refit.cuh, warp-uniform ----
    {
        bool any = false;
#pragma unroll
        for (int i = 0; i < MI; ++i)
#pragma unroll
            for (int j = 0; j < NJ; ++j)
#pragma unroll
                for (int e = 0; e < 4; ++e) {
                    tot[i][j][e] += acc[i][j][e];
                    acc[i][j][e] = 0.f;
                    any |= nonfinite32(tot[i][j][e]);
                }
        if (__any_sync(0xffffffffu, any)) {
#pragma unroll
            for (int i = 0; i < MI; ++i)
#pragma unroll
                for (int j = 0; j < NJ; ++j)
#pragma unroll
                    for (int e = 0; e < 4; ++e) {
                        const int64_t r = m0 + wm * MI * 16 + i * 16 + (e >> 1) * 8 + g;
                        const int64_t c = n0 + wn * NJ * 8 + j * 8 + 2 * t + (e & 1);
                        tot[i][j][e] = refit_real(slice_refit(p.refit, kb_begin * BK, kblocks * BK, 1), tot[i][j][e], r, c, r < p.m && c < p.n, batch);
                    }
        }
    }
    // ---- epilogue: C = alpha * (tot + acc), never read C ----
    const bool vec = ((p.n | p.ldc) & 1) == 0;
    const float alr = p.alpha.get();
#pragma unroll
    for (int i = 0; i < MI; ++i)
#pragma unroll
        for (int hr = 0; hr < 2; ++hr) {
            const int64_t r = m0 + wm * MI * 16 + i * 16 + hr * 8 + g;
            if (r >= p.m) continue;
            float *crow = p.C + batch * p.stride_c + r * p.ldc;
#pragma unroll
            for (int j = 0; j < NJ; ++j) {
                const int64_t c = n0 + wn * NJ * 8 + j * 8 + 2 * t;
                float v0 = alr * (tot[i][j][2 * hr] + acc[i][j][2 * hr]);
                float v1 = alr * (tot[i][j][2 * hr + 1] + acc[i][j][2 * hr + 1]);
                if (p.beta.on()) {
                    if (c < p.n) v0 = p.beta.add(v0, r, c);
                    if (c + 1 < p.n) v1 = p.beta.add(v1, r, c + 1);
                }
                if (vec) {
                    if (c < p.n) *reinterpret_cast<float2 *>(crow + c) = make_float2(v0, v1);
                } else {
                    if (c < p.n) crow[c] = v0;
                    if (c + 1 < p.n) crow[c + 1] = v1;
                }
            }
        }
}

// =====================================================================================
// complex f32 (interleaved).  K-block = 16 complex (128 bytes).  Fragment elements are 8-byte
// (re, im) pairs, so the shared-memory addressing is gemm_dmma.cu's: slot t of step s reads
// complex k = kslot(s, t), conflict-free per half-warp for both majors.  BM, BN in complex elements.
// =====================================================================================
using CTile64 = STile<2, 2, 2, 4, 4, 1>;    // 64 x 64 complex, 8 warps of 32 x 16
using CTile32 = STile<1, 2, 2, 2, 4, 5>;    // 32 x 32 complex, 4 warps of 16 x 16

__device__ __forceinline__ int kslot16(int s, int t) {       // = kslot() of gemm_dmma.cu
    return (t & 2) * 4 + 2 * ((s + t) & 3) + (t & 1);
}

template <bool AK, bool BKM, bool CONJA, bool CONJB, typename TL>
__global__ void __launch_bounds__(TL::THREADS, TL::MIN_CTAS)
cgemm_mma_kernel(const __grid_constant__ CUtensorMap mapA, const __grid_constant__ CUtensorMap mapB,
                 const SgemmMmaParams p)
{
    constexpr int MI = TL::MI, NJ = TL::NJ, BM = TL::BM, BN = TL::BN, STAGES = TL::STAGES, BK = 16,
                  A_BYTES = TL::A_BYTES, STAGE_BYTES = TL::STAGE_BYTES, CONSUMER_WARPS = TL::CONSUMER_WARPS;
    static_assert(NJ % 2 == 0, "MN-major B: 8-column MMA tiles pair up inside 16-wide boxes");
    extern __shared__ uint8_t smem_raw[];
    const uint32_t smem_base = (tma::smem_u32(smem_raw) + 1023u) & ~1023u;
    const uint32_t bar_base = smem_base + STAGES * STAGE_BYTES;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    int tile_m, tile_n, pa, pb;
    raster(p, tile_m, tile_n, pa, pb);
    const int m0 = tile_m * BM, n0 = tile_n * BN;
    int kb_begin = 0, kblocks = (int)((p.k + BK - 1) / BK);
    if (p.kb_per_slice > 0) {                 // split-K: the batch index is the K slice; operands are the same 2-D arrays
        kb_begin = (int)(blockIdx.x / p.tiles_per_batch) * p.kb_per_slice;
        kblocks = max(0, min(p.kb_per_slice, kblocks - kb_begin));
        pa = pb = -1;
    }

    if (threadIdx.x == 0) {
        for (int s = 0; s < STAGES; ++s) {
            tma::mbar_init(bar_base + 8 * s, 1);
            tma::mbar_init(bar_base + 8 * (STAGES + s), CONSUMER_WARPS);
        }
        tma::fence_barrier_init();
    }
    __syncthreads();

    if (warp == CONSUMER_WARPS) {
        if (lane == 0) {
            tma::prefetch_tmap(&mapA);
            tma::prefetch_tmap(&mapB);
            for (int kb = 0; kb < kblocks; ++kb) {
                const int s = kb % STAGES;
                if (kb >= STAGES) tma::mbar_wait(bar_base + 8 * (STAGES + s), ((kb / STAGES) - 1) & 1);
                const uint32_t full = bar_base + 8 * s;
                const uint32_t sa = smem_base + s * STAGE_BYTES, sb = sa + A_BYTES;
                tma::mbar_expect_tx(full, STAGE_BYTES);
                const int k0 = (kb_begin + kb) * BK;
                // tensor maps are over floats: inner coordinates are 2 x the complex index
                if (AK) tma::load_box(sa, &mapA, full, 2 * k0, m0, pa);
                else
#pragma unroll
                    for (int j = 0; j < BM / 16; ++j) tma::load_box(sa + j * 2048, &mapA, full, 2 * (m0 + 16 * j), k0, pa);
                if (BKM) tma::load_box(sb, &mapB, full, 2 * k0, n0, pb);
                else
#pragma unroll
                    for (int j = 0; j < BN / 16; ++j) tma::load_box(sb + j * 2048, &mapB, full, 2 * (n0 + 16 * j), k0, pb);
            }
        }
        return;
    }

    const int g = lane >> 2, t = lane & 3;
    const int wm = warp / TL::WN, wn = warp % TL::WN;
    uint32_t offA[4], offB[4];
#pragma unroll
    for (int s = 0; s < 4; ++s) {
        const int kk = kslot16(s, t);
        const uint32_t kmaj = g * 128 + ((((kk >> 1) ^ g) & 7) << 4) + (kk & 1) * 8;
        const uint32_t mnmaj = kk * 128 + ((((g >> 1) ^ (kk & 7)) & 7) << 4) + (g & 1) * 8;
        offA[s] = AK ? (uint32_t)(wm * MI * 16 * 128) + kmaj : (uint32_t)(wm * MI * 2048) + mnmaj;
        offB[s] = A_BYTES + (BKM ? (uint32_t)(wn * NJ * 8 * 128) + kmaj : (uint32_t)(wn * (NJ / 2) * 2048) + mnmaj);
    }

    float are[MI][NJ][4], aim[MI][NJ][4], tre[MI][NJ][4], tim[MI][NJ][4];
#pragma unroll
    for (int i = 0; i < MI; ++i)
#pragma unroll
        for (int j = 0; j < NJ; ++j)
#pragma unroll
            for (int e = 0; e < 4; ++e) are[i][j][e] = aim[i][j][e] = tre[i][j][e] = tim[i][j][e] = 0.f;

    for (int kb = 0; kb < kblocks; ++kb) {
        const int s = kb % STAGES;
        const uint32_t stage = smem_base + s * STAGE_BYTES;
        tma::mbar_wait(bar_base + 8 * s, (kb / STAGES) & 1);
#pragma unroll
        for (int st = 0; st < 4; ++st) {
            uint32_t ahi[MI][4], alo[MI][4];
            uint32_t behi[NJ][2], belo[NJ][2], bohi[NJ][2], bolo[NJ][2];   // "even" (br,-bi) and "odd" (bi,br) fragments
#pragma unroll
            for (int i = 0; i < MI; ++i) {
                const uint32_t lo_rows = AK ? stage + offA[st] + i * 2048 : stage + offA[st] + i * 2048;
                const uint32_t hi_rows = AK ? lo_rows + 1024 : stage + (offA[st] ^ 64u) + i * 2048;
                const float2 v0 = lds64f(lo_rows), v1 = lds64f(hi_rows);
                split_tf32(v0.x, ahi[i][0], alo[i][0]);                    // a0 = Re A[g][l]     (embedded k = 2l)
                split_tf32(v1.x, ahi[i][1], alo[i][1]);                    // a1 = Re A[g+8][l]
                split_tf32(v0.y, ahi[i][2], alo[i][2]);                    // a2 = Im A[g][l]     (embedded k = 2l+1)
                split_tf32(v1.y, ahi[i][3], alo[i][3]);
                if (CONJA) { ahi[i][2] ^= 0x80000000u; alo[i][2] ^= 0x80000000u; ahi[i][3] ^= 0x80000000u; alo[i][3] ^= 0x80000000u; }
            }
#pragma unroll
            for (int j = 0; j < NJ; ++j) {
                const uint32_t addr = BKM ? stage + offB[st] + j * 1024
                                          : stage + (offB[st] ^ (uint32_t)((j & 1) << 6)) + (j >> 1) * 2048;
                const float2 v = lds64f(addr);
                uint32_t rh, rl, ih, il;
                split_tf32(v.x, rh, rl);
                split_tf32(v.y, ih, il);
                if (CONJB) { ih ^= 0x80000000u; il ^= 0x80000000u; }
                behi[j][0] = rh; belo[j][0] = rl; behi[j][1] = ih ^ 0x80000000u; belo[j][1] = il ^ 0x80000000u;
                bohi[j][0] = ih; bolo[j][0] = il; bohi[j][1] = rh; bolo[j][1] = rl;
            }
#pragma unroll
            for (int i = 0; i < MI; ++i)
#pragma unroll
                for (int j = 0; j < NJ; ++j) {
                    mma_3x(are[i][j], ahi[i], alo[i], behi[j], belo[j]);
                    mma_3x(aim[i][j], ahi[i], alo[i], bohi[j], bolo[j]);
                }
        }
        tma::consumer_release_fence();
        if (lane == 0) tma::mbar_arrive(bar_base + 8 * (STAGES + s));
        if ((kb + 1) % PROMO_KB == 0) {
#pragma unroll
            for (int i = 0; i < MI; ++i)
#pragma unroll
                for (int j = 0; j < NJ; ++j)
#pragma unroll
                    for (int e = 0; e < 4; ++e) {
                        tre[i][j][e] += are[i][j][e]; are[i][j][e] = 0.f;
                        tim[i][j][e] += aim[i][j][e]; aim[i][j][e] = 0.f;
                    }
        }
    }

    const int64_t batch = p.tiles_per_batch > 0 ? blockIdx.x / p.tiles_per_batch : 0;
    // ---- non-finite sums: refit.cuh, warp-uniform ----
    {
        bool any = false;
#pragma unroll
        for (int i = 0; i < MI; ++i)
#pragma unroll
            for (int j = 0; j < NJ; ++j)
#pragma unroll
                for (int e = 0; e < 4; ++e) {
                    tre[i][j][e] += are[i][j][e]; are[i][j][e] = 0.f;
                    tim[i][j][e] += aim[i][j][e]; aim[i][j][e] = 0.f;
                    any |= nonfinite32(tre[i][j][e]) || nonfinite32(tim[i][j][e]);
                }
        if (__any_sync(0xffffffffu, any)) {
#pragma unroll
            for (int i = 0; i < MI; ++i)
#pragma unroll
                for (int j = 0; j < NJ; ++j)
#pragma unroll
                    for (int e = 0; e < 4; ++e) {
                        const int64_t r = m0 + wm * MI * 16 + i * 16 + (e >> 1) * 8 + g;
                        const int64_t c = n0 + wn * NJ * 8 + j * 8 + 2 * t + (e & 1);
                        const float2 v = refit_cplx(slice_refit(p.refit, kb_begin * BK, kblocks * BK, 2), make_float2(tre[i][j][e], tim[i][j][e]), r, c, r < p.m && c < p.n, batch);
                        tre[i][j][e] = v.x; tim[i][j][e] = v.y;
                    }
        }
    }
    float2 *C = reinterpret_cast<float2 *>(p.C) + batch * p.stride_c;
    const float2 al = p.calpha.get();
#pragma unroll
    for (int i = 0; i < MI; ++i)
#pragma unroll
        for (int hr = 0; hr < 2; ++hr) {
            const int64_t r = m0 + wm * MI * 16 + i * 16 + hr * 8 + g;
            if (r >= p.m) continue;
            float2 *crow = C + r * p.ldc;
#pragma unroll
            for (int j = 0; j < NJ; ++j)
#pragma unroll
                for (int e = 0; e < 2; ++e) {
                    const int64_t c = n0 + wn * NJ * 8 + j * 8 + 2 * t + e;
                    if (c < p.n) {
                        const float xr = tre[i][j][2 * hr + e] + are[i][j][2 * hr + e];
                        const float xi = tim[i][j][2 * hr + e] + aim[i][j][2 * hr + e];
                        float2 v = make_float2(al.x * xr - al.y * xi, al.x * xi + al.y * xr);
                        if (p.cbeta.on()) v = p.cbeta.add(v, r, c);
                        crow[c] = v;
                    }
                }
        }
}

// ---------------------------------------------------------------------------------
// host side
// ---------------------------------------------------------------------------------
bool aligned16(const void *p) { return (reinterpret_cast<uintptr_t>(p) & 15u) == 0; }

// tensor maps of one product (2-D) or of a strided batch (3-D, tma::make_2d_batched); `fpe` = floats per element
bool mma_maps(CUtensorMap *mapA, CUtensorMap *mapB, const GemmArgs &g, int fpe, uint32_t box_rows_a, uint32_t box_rows_b)
{
    const CUtensorMapDataType F = CU_TENSOR_MAP_DATA_TYPE_FLOAT32;
    const uint64_t eb = 4 * (uint64_t)fpe;
    if (g.batch > 1)
        return tma::make_2d_batched(mapA, F, 4, g.A, (uint64_t)g.rowsA, fpe * (uint64_t)g.colsA, box_rows_a, 32, (uint64_t)g.lda * eb,
                                    g.strideA ? (uint64_t)g.batch : 1, (uint64_t)(g.strideA ? g.strideA : g.rowsA * g.lda) * eb) &&
               tma::make_2d_batched(mapB, F, 4, g.B, (uint64_t)g.rowsB, fpe * (uint64_t)g.colsB, box_rows_b, 32, (uint64_t)g.ldb * eb,
                                    g.strideB ? (uint64_t)g.batch : 1, (uint64_t)(g.strideB ? g.strideB : g.rowsB * g.ldb) * eb);
    return tma::make_2d(mapA, F, 4, g.A, (uint64_t)g.rowsA, fpe * (uint64_t)g.colsA, box_rows_a, 32, (uint64_t)g.lda * eb) &&
           tma::make_2d(mapB, F, 4, g.B, (uint64_t)g.rowsB, fpe * (uint64_t)g.colsB, box_rows_b, 32, (uint64_t)g.ldb * eb);
}
// grid size: tiles of one product, times the batch
int64_t mma_batch_grid(SgemmMmaParams &p, const GemmArgs &g)
{
    int64_t tiles = (int64_t)p.tiles_m * p.tiles_n;
    p.tiles_per_batch = 0; p.plane_a = p.plane_b = 0; p.stride_c = 0; p.kb_per_slice = 0;
    if (g.ksplit > 1 && g.batch <= 1) {              // split-K: slices as the batch index, partials into the scratch planes
        const int bk = g.dtype == ACCBLAS_C32 ? 16 : 32;
        const int kblocks = (int)((g.k + bk - 1) / bk);
        p.tiles_per_batch = (int)tiles;
        p.kb_per_slice = (kblocks + g.ksplit - 1) / g.ksplit;
        p.C = static_cast<float *>(g.ksplit_ws); p.ldc = g.n; p.stride_c = g.m * g.n;
        return tiles * g.ksplit;
    }
    if (g.batch > 1) {
        p.tiles_per_batch = (int)tiles;
        p.plane_a = g.strideA ? 1 : 0; p.plane_b = g.strideB ? 1 : 0;
        p.stride_c = g.strideC;
        tiles *= g.batch;
    }
    return tiles;
}

template <bool AK, bool BKM, typename TL>
int launch_sgemm_mma(const GemmArgs &g)
{
    constexpr int BM = TL::BM, BN = TL::BN;
    CUtensorMap mapA, mapB;
    // K-major operand: box = 32 (k) x BM|BN rows; MN-major operand: box = 32 (m|n) x 32 (k)
    if (!mma_maps(&mapA, &mapB, g, 1, AK ? BM : 32, BKM ? BN : 32)) return cuda_fail(cudaErrorInvalidValue, "cuTensorMapEncodeTiled");
    static thread_local AttrOnce attr_once;
    if (attr_once.need()) {
        ACC_CUDA(cudaFuncSetAttribute(sgemm_mma_kernel<AK, BKM, TL>, cudaFuncAttributeMaxDynamicSharedMemorySize, TL::SMEM_BYTES));
    }
    SgemmMmaParams p;
    p.m = g.m; p.n = g.n; p.k = g.k; p.ldc = g.ldc;
    p.alpha = make_alpha<float>(g.alpha, g.alpha_dev); p.calpha = Alpha<float2>{make_float2(0.f, 0.f), nullptr};
    p.beta = make_beta<float>(g); p.cbeta = BetaC<float2>{nullptr, 0, make_float2(0.f, 0.f)};
    p.C = static_cast<float *>(g.C);
    p.refit = make_refit32(g);
    p.tiles_m = (int)((g.m + BM - 1) / BM);
    p.tiles_n = (int)((g.n + BN - 1) / BN);
    const int64_t tiles = mma_batch_grid(p, g);
    if (tiles > 0x7fffffff) return ACCBLAS_ERR_SHAPE;
    sgemm_mma_kernel<AK, BKM, TL><<<(unsigned)tiles, TL::THREADS, TL::SMEM_BYTES, g.stream>>>(mapA, mapB, p);
    return check_launch("sgemm_mma_kernel");
}

// DIRECT launch (operands that miss TMA's alignment rules, whole K resident in shared memory): 32 x 32 tiles, one stage per
// K-block.  The tensor-map parameters are unused placeholders.
constexpr int kDirectMaxKBlocks = 24;            // 24 x 8 KiB stages + slack fit the 227 KiB a CTA may use (k <= 768)
template <bool AK, bool BKM>
int launch_sgemm_mma_direct(const GemmArgs &g)
{
    using TL = STile32;
    constexpr int MAX_SMEM = kDirectMaxKBlocks * TL::STAGE_BYTES + 1024;
    static thread_local AttrOnce attr_once;
    if (attr_once.need()) {
        ACC_CUDA(cudaFuncSetAttribute(sgemm_mma_kernel<AK, BKM, TL, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, MAX_SMEM));
    }
    SgemmMmaParams p;
    p.m = g.m; p.n = g.n; p.k = g.k; p.ldc = g.ldc;
    p.alpha = make_alpha<float>(g.alpha, g.alpha_dev); p.calpha = Alpha<float2>{make_float2(0.f, 0.f), nullptr};
    p.beta = make_beta<float>(g); p.cbeta = BetaC<float2>{nullptr, 0, make_float2(0.f, 0.f)};
    p.C = static_cast<float *>(g.C);
    p.refit = make_refit32(g);                    // also carries the operand pointers and pitches the loader reads
    p.tiles_m = (int)((g.m + TL::BM - 1) / TL::BM);
    p.tiles_n = (int)((g.n + TL::BN - 1) / TL::BN);
    p.tiles_per_batch = 0; p.plane_a = p.plane_b = 0; p.stride_c = 0; p.kb_per_slice = 0;
    const int64_t tiles = (int64_t)p.tiles_m * p.tiles_n;
    if (tiles > 0x7fffffff) return ACCBLAS_ERR_SHAPE;
    const int kblocks = (int)((g.k + 31) / 32);
    static const CUtensorMap unused{};
    sgemm_mma_kernel<AK, BKM, TL, true><<<(unsigned)tiles, kDirectThreads, (size_t)kblocks * TL::STAGE_BYTES + 1024, g.stream>>>(unused, unused, p);
    return check_launch("sgemm_mma_kernel<direct>");
}

template <bool AK, bool BKM, bool CA, bool CB, typename TL>
int launch_cgemm_mma(const GemmArgs &g)
{
    constexpr int BM = TL::BM, BN = TL::BN;
    CUtensorMap mapA, mapB;
    // maps are over floats (a complex row of c columns is 2c floats); boxes are 32 floats = 16 complex wide
    if (!mma_maps(&mapA, &mapB, g, 2, AK ? BM : 16, BKM ? BN : 16)) return cuda_fail(cudaErrorInvalidValue, "cuTensorMapEncodeTiled");
    static thread_local AttrOnce attr_once;
    if (attr_once.need()) {
        ACC_CUDA(cudaFuncSetAttribute(cgemm_mma_kernel<AK, BKM, CA, CB, TL>, cudaFuncAttributeMaxDynamicSharedMemorySize, TL::SMEM_BYTES));
    }
    SgemmMmaParams p;
    p.m = g.m; p.n = g.n; p.k = g.k; p.ldc = g.ldc;
    p.calpha = make_alpha<float2>(g.alpha, g.alpha_dev); p.alpha = Alpha<float>{0.f, nullptr};
    p.cbeta = make_beta<float2>(g); p.beta = BetaC<float>{nullptr, 0, 0.f};
    p.C = static_cast<float *>(g.C);
    p.refit = make_refit32(g);
    p.tiles_m = (int)((g.m + BM - 1) / BM);
    p.tiles_n = (int)((g.n + BN - 1) / BN);
    const int64_t tiles = mma_batch_grid(p, g);
    if (tiles > 0x7fffffff) return ACCBLAS_ERR_SHAPE;
    cgemm_mma_kernel<AK, BKM, CA, CB, TL><<<(unsigned)tiles, TL::THREADS, TL::SMEM_BYTES, g.stream>>>(mapA, mapB, p);
    return check_launch("cgemm_mma_kernel");
}

// busiest-SM cost of a tiling (see pick_tile in gemm_dmma.cu)
struct TileModel { int bm, bn; double eff, eff_alone; };
// efficiencies fitted to profiles/r01_sgemm_families.jsonl (64 x 64 is the best real tile at every size it fills)
constexpr TileModel kSTiles[3] = {{128, 64, 0.85, 0.85}, {64, 64, 1.00, 1.00}, {32, 32, 0.66, 0.48}};
constexpr TileModel kCTiles[2] = {{64, 64, 1.00, 1.00}, {32, 32, 0.80, 0.50}};

int pick(const TileModel *tm, int count, int forced, int64_t m, int64_t n, int64_t batch = 1)
{
    if (forced >= 0 && forced < count) return forced;
    const int sms = sm_count();
    int best = 0;
    double best_cost = 0.0;
    for (int c = 0; c < count; ++c) {
        const int64_t tiles = ((m + tm[c].bm - 1) / tm[c].bm) * ((n + tm[c].bn - 1) / tm[c].bn) * batch;
        const double per_sm = (double)((tiles + sms - 1) / sms);
        const double cost = per_sm * tm[c].bm * tm[c].bn / (per_sm > 1.0 ? tm[c].eff : tm[c].eff_alone);
        if (c == 0 || cost < best_cost * 0.999) { best = c; best_cost = cost; }
    }
    return best;
}

// Split-K over idle SMs for the 32 x 32 tiles (as in gemm_dmma.cu): at most half a wave of tiles, a long K -> up to 16 K
// slices as the batch index of ONE launch, alpha-scaled partials into scratch, summed in slice order by splitk_reduce.
int splitk_slices(const GemmArgs &g, int bk)
{
    if (g.batch > 1 || has_beta(g) || g.ksplit != 0 || g_tuning[ACCBLAS_TUNE_SPLITK] == 0) return 0;
    const int64_t tiles32 = ((g.m + 31) / 32) * ((g.n + 31) / 32), kblocks = (g.k + bk - 1) / bk;
    // a K-block costs a one-wave CTA ~0.45 us (real, 32 wide) / ~0.9 us (complex, 16 wide); the second launch ~6 us in all:
    // split from 16 (8) K-blocks of saving, in slices of at least 4
    const int64_t need = g.dtype == ACCBLAS_C32 ? 8 : 16;
    if (2 * tiles32 > sm_count() || kblocks < 12) return 0;
    const int S = (int)std::min<int64_t>(std::min<int64_t>(2 * sm_count() / tiles32, kblocks / 4), 16);
    return (S >= 2 && kblocks * (S - 1) >= need * S) ? S : 0;
}
template <typename Launch>
int with_splitk(const GemmArgs &g, int S, Launch launch)
{
    GemmArgs h = g;
    h.ksplit = S;
    ACC_CUDA(scratch_alloc(&h.ksplit_ws, elem_size(g.dtype) * (size_t)S * (size_t)g.m * (size_t)g.n, g.stream));
    int st = launch(h);
    if (st == ACCBLAS_OK) st = splitk_reduce(h);
    const cudaError_t e = cudaFreeAsync(h.ksplit_ws, g.stream);
    if (st == ACCBLAS_OK && e != cudaSuccess) return cuda_fail(e, "cudaFreeAsync");
    return st;
}

template <bool AK, bool BKM>
int sgemm_mma_tiles(const GemmArgs &g)
{
    switch (pick(kSTiles, 3, g_tuning[ACCBLAS_TUNE_SMMA_TILE], g.m, g.n, g.batch > 1 ? g.batch : 1)) {
    case 1:  return launch_sgemm_mma<AK, BKM, STile64>(g);
    case 2:  {
        const int S = splitk_slices(g, 32);
        if (S >= 2) return with_splitk(g, S, [](const GemmArgs &h) { return launch_sgemm_mma<AK, BKM, STile32>(h); });
        return launch_sgemm_mma<AK, BKM, STile32>(g);
    }
    default: return launch_sgemm_mma<AK, BKM, STile128>(g);
    }
}

template <bool AK, bool BKM, bool CA, bool CB>
int cgemm_mma_tiles(const GemmArgs &g)
{
    const int forced = g_tuning[ACCBLAS_TUNE_SMMA_TILE] < 0 ? -1 : (g_tuning[ACCBLAS_TUNE_SMMA_TILE] >= 2 ? 1 : 0);
    if (pick(kCTiles, 2, forced, g.m, g.n, g.batch > 1 ? g.batch : 1) == 1) {
        const int S = splitk_slices(g, 16);
        if (S >= 2) return with_splitk(g, S, [](const GemmArgs &h) { return launch_cgemm_mma<AK, BKM, CA, CB, CTile32>(h); });
        return launch_cgemm_mma<AK, BKM, CA, CB, CTile32>(g);
    }
    return launch_cgemm_mma<AK, BKM, CA, CB, CTile64>(g);
}

template <bool AK, bool BKM>
int cgemm_mma_conj(const GemmArgs &g)
{
    // conjugation only ever comes with a transpose (op H): CA implies !AK, CB implies BKM
    const bool ca = g.opA == ACCBLAS_OP_H, cb = g.opB == ACCBLAS_OP_H;
    if constexpr (!AK && BKM) {
        if (ca && cb) return cgemm_mma_tiles<AK, BKM, true, true>(g);
        if (ca) return cgemm_mma_tiles<AK, BKM, true, false>(g);
        if (cb) return cgemm_mma_tiles<AK, BKM, false, true>(g);
    } else if constexpr (!AK) {
        if (ca) return cgemm_mma_tiles<AK, BKM, true, false>(g);
    } else if constexpr (BKM) {
        if (cb) return cgemm_mma_tiles<AK, BKM, false, true>(g);
    }
    return cgemm_mma_tiles<AK, BKM, false, false>(g);
}

}  // namespace

bool sgemm_mma_eligible(const GemmArgs &g)
{
    if (g.dtype != ACCBLAS_F32 && g.dtype != ACCBLAS_C32) return false;
    if (g.m < 1 || g.n < 1 || g.k < 1) return false;
    if (!aligned16(g.A) || !aligned16(g.B) || !aligned16(g.C)) return false;
    const int64_t lim = g.dtype == ACCBLAS_C32 ? 0x3fffffff : 0x7fffffff;
    if (g.rowsA > lim || g.colsA > lim || g.rowsB > lim || g.colsB > lim) return false;
    // TMA: row pitch a multiple of 16 bytes
    const int64_t q = g.dtype == ACCBLAS_C32 ? 2 : 4;
    if ((g.lda % q) || (g.ldb % q)) return false;
    // tiny products stay on the SIMT kernel: a 32 x 32 tile would be mostly padding
    return g.m * g.n >= 24 * 24 && g.k >= 4;
}

// f32 products with a misaligned operand (base or row pitch off TMA's 16-byte grid) whose K extent fits in shared memory
bool sgemm_mma_direct_eligible(const GemmArgs &g)
{
    if (g.dtype != ACCBLAS_F32 || g.batch > 1) return false;
    if (!aligned16(g.C) || g.m * g.n < 24 * 24 || g.k < 4) return false;
    if (g.k > 32 * kDirectMaxKBlocks) return false;
    return g.rowsA <= 0x7fffffff && g.colsA <= 0x7fffffff && g.rowsB <= 0x7fffffff && g.colsB <= 0x7fffffff;
}

int sgemm_mma_direct(const GemmArgs &g)
{
    const bool ak = g.opA == ACCBLAS_OP_N, bk = g.opB != ACCBLAS_OP_N;
    if (ak && bk) return launch_sgemm_mma_direct<true, true>(g);
    if (ak && !bk) return launch_sgemm_mma_direct<true, false>(g);
    if (!ak && bk) return launch_sgemm_mma_direct<false, true>(g);
    return launch_sgemm_mma_direct<false, false>(g);
}

int sgemm_mma(const GemmArgs &g)
{
    const bool ak = g.opA == ACCBLAS_OP_N, bk = g.opB != ACCBLAS_OP_N;
    if (g.dtype == ACCBLAS_C32) {
        if (ak && bk) return cgemm_mma_conj<true, true>(g);
        if (ak && !bk) return cgemm_mma_conj<true, false>(g);
        if (!ak && bk) return cgemm_mma_conj<false, true>(g);
        return cgemm_mma_conj<false, false>(g);
    }
    if (ak && bk) return sgemm_mma_tiles<true, true>(g);
    if (ak && !bk) return sgemm_mma_tiles<true, false>(g);
    if (!ak && bk) return sgemm_mma_tiles<false, true>(g);
    return sgemm_mma_tiles<false, false>(g);
}

}  // namespace accblas
