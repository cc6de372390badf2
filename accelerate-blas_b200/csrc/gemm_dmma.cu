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

namespace accblas {
namespace {

constexpr int BN = 128, BK = 16;
constexpr int CONSUMER_WARPS = 8;
constexpr int THREADS = (CONSUMER_WARPS + 1) * 32;
constexpr int B_BYTES = BN * BK * 8;
// Two CTA tiles: MI = 8 -> 128 x 128 (one CTA per SM, 6 stages), the throughput shape; MI = 4 -> 64 x 128
// (two CTAs per SM, 4 stages each) for products whose 128 x 128 tiling leaves SMs idle in the last
// wave (dgemm 1024^3 is 64 big tiles on 148 SMs).  MI = 8-row DMMA tiles per warp along m.
template <int MI> struct Cfg {
    static constexpr int BM = 16 * MI;
    static constexpr int STAGES = MI == 8 ? 6 : 4;
    static constexpr int A_BYTES = BM * BK * 8;
    static constexpr int STAGE_BYTES = A_BYTES + B_BYTES;
    static constexpr int SMEM_BYTES = STAGES * STAGE_BYTES + 1024 /*align slack*/ + 2 * STAGES * 8;
    static constexpr int MIN_CTAS = MI == 8 ? 1 : 2;
};

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

struct DgemmParams {
    int64_t m, n, k, ldc;
    double alpha;
    double *C;
    int tiles_m, tiles_n;
};

// AK: op(A) is K-major in memory (opA == N).  BKM: op(B) is K-major in memory (opB == T).
template <bool AK, bool BKM, int MI>
__global__ void __launch_bounds__(THREADS, Cfg<MI>::MIN_CTAS)
dgemm_dmma_kernel(const __grid_constant__ CUtensorMap mapA, const __grid_constant__ CUtensorMap mapB,
                  const DgemmParams p)
{
    constexpr int BM = Cfg<MI>::BM, STAGES = Cfg<MI>::STAGES, A_BYTES = Cfg<MI>::A_BYTES,
                  STAGE_BYTES = Cfg<MI>::STAGE_BYTES;
    extern __shared__ uint8_t smem_raw[];
    const uint32_t smem_base = (tma::smem_u32(smem_raw) + 1023u) & ~1023u;
    const uint32_t bar_base = smem_base + STAGES * STAGE_BYTES;      // full[STAGES], empty[STAGES]
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

    // grouped rasterisation: 16 tile-rows per group so that one resident wave re-uses its
    // A and B panels out of L2
    constexpr int GROUP = 16;
    int tile_m, tile_n;
    {
        const int per_group = GROUP * p.tiles_n;
        const int grp = blockIdx.x / per_group;
        const int first_m = grp * GROUP;
        const int gsz = min(p.tiles_m - first_m, GROUP);
        const int in_grp = blockIdx.x - grp * per_group;
        tile_m = first_m + in_grp % gsz;
        tile_n = in_grp / gsz;
    }
    const int m0 = tile_m * BM, n0 = tile_n * BN;
    const int kblocks = (int)((p.k + BK - 1) / BK);

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
            for (int kb = 0; kb < kblocks; ++kb) {
                const int s = kb % STAGES;
                const uint32_t round = kb / STAGES;
                if (kb >= STAGES) tma::mbar_wait(bar_base + 8 * (STAGES + s), (round - 1) & 1);
                const uint32_t full = bar_base + 8 * s;
                const uint32_t sa = smem_base + s * STAGE_BYTES, sb = sa + A_BYTES;
                tma::mbar_expect_tx(full, STAGE_BYTES);
                const int k0 = kb * BK;
                if (AK) tma::load_2d(sa, &mapA, full, k0, m0);
                else
#pragma unroll
                    for (int j = 0; j < BM / 16; ++j) tma::load_2d(sa + j * 2048, &mapA, full, m0 + 16 * j, k0);
                if (BKM) tma::load_2d(sb, &mapB, full, k0, n0);
                else
#pragma unroll
                    for (int j = 0; j < BN / 16; ++j) tma::load_2d(sb + j * 2048, &mapB, full, n0 + 16 * j, k0);
            }
        }
        return;
    }

    // ================================ consumers ================================
    const int g = lane >> 2, t = lane & 3;
    const int wm = warp >> 2, wn = warp & 3;          // 2 x 4 warps, warp tile (8 MI) x 32
    // per-lane byte offsets of the four k-steps inside a stage tile (frag / box offsets are added
    // as immediates below)
    uint32_t offA[4], offB[4];
#pragma unroll
    for (int s = 0; s < 4; ++s) {
        const int kk = kslot(s, t);
        const uint32_t kmaj = g * 128 + ((((kk >> 1) ^ g) & 7) << 4) + (kk & 1) * 8;
        const uint32_t mnmaj = kk * 128 + ((((g >> 1) ^ (kk & 7)) & 7) << 4) + (g & 1) * 8;
        offA[s] = AK ? (uint32_t)(wm * MI * 8 * 128) + kmaj : (uint32_t)(wm * (MI / 2) * 2048) + mnmaj;
        offB[s] = A_BYTES + (BKM ? (uint32_t)(wn * 32 * 128) + kmaj : (uint32_t)(wn * 2 * 2048) + mnmaj);
    }

    double acc[MI][4][2];
#pragma unroll
    for (int i = 0; i < MI; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j][0] = acc[i][j][1] = 0.0;

    auto load_frags = [&](uint32_t stage, int s, double (&a)[MI], double (&b)[4]) {
#pragma unroll
        for (int i = 0; i < MI; ++i)
            a[i] = lds64(stage + (AK ? offA[s] + i * 1024 : (offA[s] ^ ((i & 1) << 6)) + (i >> 1) * 2048));
#pragma unroll
        for (int j = 0; j < 4; ++j)
            b[j] = lds64(stage + (BKM ? offB[s] + j * 1024 : (offB[s] ^ ((j & 1) << 6)) + (j >> 1) * 2048));
    };

    for (int kb = 0; kb < kblocks; ++kb) {
        const int s = kb % STAGES;
        const uint32_t stage = smem_base + s * STAGE_BYTES;
        tma::mbar_wait(bar_base + 8 * s, (kb / STAGES) & 1);
        double a[2][MI], b[2][4];
        load_frags(stage, 0, a[0], b[0]);
#pragma unroll
        for (int st = 0; st < 4; ++st) {
            if (st < 3) load_frags(stage, st + 1, a[(st + 1) & 1], b[(st + 1) & 1]);
#pragma unroll
            for (int i = 0; i < MI; ++i)
#pragma unroll
                for (int j = 0; j < 4; ++j) dmma(acc[i][j][0], acc[i][j][1], a[st & 1][i], b[st & 1][j]);
        }
        __syncwarp();
        if (lane == 0) tma::mbar_arrive(bar_base + 8 * (STAGES + s));   // stage consumed by this warp
    }

    // ---- epilogue: C = alpha * acc (never read C) ----
    const bool vec = ((p.n | p.ldc) & 1) == 0;
#pragma unroll
    for (int i = 0; i < MI; ++i) {
        const int64_t r = m0 + wm * MI * 8 + i * 8 + g;
        if (r >= p.m) continue;
        double *crow = p.C + r * p.ldc;
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const int64_t c = n0 + wn * 32 + j * 8 + 2 * t;
            const double v0 = p.alpha * acc[i][j][0], v1 = p.alpha * acc[i][j][1];
            if (vec) {
                if (c < p.n) *reinterpret_cast<double2 *>(crow + c) = make_double2(v0, v1);
            } else {
                if (c < p.n) crow[c] = v0;
                if (c + 1 < p.n) crow[c + 1] = v1;
            }
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
constexpr int ZBM = 128, ZBN = 64, ZBK = 8;
constexpr int ZSTAGES = 8;
constexpr int ZTHREADS = 384;   // 2 consumer warpgroups + 1 producer warpgroup (setmaxnreg is per warpgroup)
constexpr int ZA_BYTES = ZBM * ZBK * 16, ZB_BYTES = ZBN * ZBK * 16, ZSTAGE_BYTES = ZA_BYTES + ZB_BYTES;
constexpr int ZSMEM_BYTES = ZSTAGES * ZSTAGE_BYTES + 1024 + 2 * ZSTAGES * 8;

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
    double alr, ali;
    double2 *C;
    int tiles_m, tiles_n;
};

template <bool AK, bool BKM, bool CONJA, bool CONJB>
__global__ void __launch_bounds__(ZTHREADS, 1)
zgemm_dmma_kernel(const __grid_constant__ CUtensorMap mapA, const __grid_constant__ CUtensorMap mapB,
                  const ZgemmParams p)
{
    extern __shared__ uint8_t smem_raw[];
    const uint32_t smem_base = (tma::smem_u32(smem_raw) + 1023u) & ~1023u;
    const uint32_t bar_base = smem_base + ZSTAGES * ZSTAGE_BYTES;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

    constexpr int GROUP = 16;
    int tile_m, tile_n;
    {
        const int per_group = GROUP * p.tiles_n;
        const int grp = blockIdx.x / per_group;
        const int first_m = grp * GROUP;
        const int gsz = min(p.tiles_m - first_m, GROUP);
        const int in_grp = blockIdx.x - grp * per_group;
        tile_m = first_m + in_grp % gsz;
        tile_n = in_grp / gsz;
    }
    const int m0 = tile_m * ZBM, n0 = tile_n * ZBN;
    const int kblocks = (int)((p.k + ZBK - 1) / ZBK);

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
        asm volatile("setmaxnreg.dec.sync.aligned.u32 40;");
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
                const int k0 = kb * ZBK;
                // tensor maps are over doubles: inner coordinates are 2 x the complex index
                if (AK) tma::load_2d(sa, &mapA, full, 2 * k0, m0);
                else
#pragma unroll
                    for (int j = 0; j < ZBM / 8; ++j) tma::load_2d(sa + j * 1024, &mapA, full, 2 * (m0 + 8 * j), k0);
                if (BKM) tma::load_2d(sb, &mapB, full, 2 * k0, n0);
                else
#pragma unroll
                    for (int j = 0; j < ZBN / 8; ++j) tma::load_2d(sb + j * 1024, &mapB, full, 2 * (n0 + 8 * j), k0);
            }
        }
        return;
    }

    asm volatile("setmaxnreg.inc.sync.aligned.u32 232;");
    const int g = lane >> 2, t = lane & 3;
    const int wm = warp >> 1, wn = warp & 1;          // 4 x 2 warps, warp tile 32 x 32 complex
    uint32_t offA[2], offB[2];
#pragma unroll
    for (int s = 0; s < 2; ++s) {
        const int kk = 2 * t + s;
        const uint32_t kmaj = g * 128 + (((kk ^ g) & 7) << 4);
        const uint32_t mnmaj = kk * 128 + (((g ^ kk) & 7) << 4);
        offA[s] = (uint32_t)(wm * 4 * 1024) + (AK ? kmaj : mnmaj);
        offB[s] = ZA_BYTES + (uint32_t)(wn * 4 * 1024) + (BKM ? kmaj : mnmaj);
    }

    double cr[4][4][2], ci[4][4][2];
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) cr[i][j][0] = cr[i][j][1] = ci[i][j][0] = ci[i][j][1] = 0.0;

    for (int kb = 0; kb < kblocks; ++kb) {
        const int s = kb % ZSTAGES;
        const uint32_t stage = smem_base + s * ZSTAGE_BYTES;
        tma::mbar_wait(bar_base + 8 * s, (kb / ZSTAGES) & 1);
        double2 a[2][4], b[2][4];
#pragma unroll
        for (int i = 0; i < 4; ++i) a[0][i] = lds128(stage + offA[0] + i * 1024);
#pragma unroll
        for (int j = 0; j < 4; ++j) b[0][j] = lds128(stage + offB[0] + j * 1024);
#pragma unroll
        for (int st = 0; st < 2; ++st) {
            if (st == 0) {
#pragma unroll
                for (int i = 0; i < 4; ++i) a[1][i] = lds128(stage + offA[1] + i * 1024);
#pragma unroll
                for (int j = 0; j < 4; ++j) b[1][j] = lds128(stage + offB[1] + j * 1024);
            }
            double ai[4], nai[4], bi[4];
#pragma unroll
            for (int i = 0; i < 4; ++i) {       // ai = Im(op A), nai = -Im(op A): one sign flip either way
                ai[i] = CONJA ? flip(a[st][i].y) : a[st][i].y;
                nai[i] = CONJA ? a[st][i].y : flip(a[st][i].y);
            }
#pragma unroll
            for (int j = 0; j < 4; ++j) bi[j] = CONJB ? flip(b[st][j].y) : b[st][j].y;
            // four passes over the 4 x 4 tile grid: 16 independent DMMAs between dependent ones
#pragma unroll
            for (int i = 0; i < 4; ++i)
#pragma unroll
                for (int j = 0; j < 4; ++j) dmma(cr[i][j][0], cr[i][j][1], a[st][i].x, b[st][j].x);
#pragma unroll
            for (int i = 0; i < 4; ++i)
#pragma unroll
                for (int j = 0; j < 4; ++j) dmma(ci[i][j][0], ci[i][j][1], a[st][i].x, bi[j]);
#pragma unroll
            for (int i = 0; i < 4; ++i)
#pragma unroll
                for (int j = 0; j < 4; ++j) dmma(cr[i][j][0], cr[i][j][1], nai[i], bi[j]);
#pragma unroll
            for (int i = 0; i < 4; ++i)
#pragma unroll
                for (int j = 0; j < 4; ++j) dmma(ci[i][j][0], ci[i][j][1], ai[i], b[st][j].x);
        }
        __syncwarp();
        if (lane == 0) tma::mbar_arrive(bar_base + 8 * (ZSTAGES + s));
    }

#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const int64_t r = m0 + wm * 32 + i * 8 + g;
        if (r >= p.m) continue;
        double2 *crow = p.C + r * p.ldc;
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const int64_t c = n0 + wn * 32 + j * 8 + 2 * t;
#pragma unroll
            for (int e = 0; e < 2; ++e) {
                if (c + e < p.n) {
                    const double xr = cr[i][j][e], xi = ci[i][j][e];
                    crow[c + e] = make_double2(p.alr * xr - p.ali * xi, p.alr * xi + p.ali * xr);
                }
            }
        }
    }
}

// ---------------------------------------------------------------------------------
// host side (real f64)
// ---------------------------------------------------------------------------------
bool aligned16(const void *p) { return (reinterpret_cast<uintptr_t>(p) & 15u) == 0; }

// Time of a tiling in units of "one full wave of 128 x 128 tiles": two co-resident 64 x 128 CTAs share
// the SM's DMMA pipe, so a wave of 296 half tiles costs the same as a wave of 148 full ones, and a
// last wave of at most one CTA per SM costs half.
bool prefer_small_tile(int64_t m, int64_t n)
{
    const int sms = sm_count();
    const int64_t t128 = ((m + 127) / 128) * ((n + BN - 1) / BN);
    const int64_t t64 = ((m + 63) / 64) * ((n + BN - 1) / BN);
    const double big = (double)((t128 + sms - 1) / sms);
    const int64_t rem = t64 % (2 * sms);
    const double small = (double)(t64 / (2 * sms)) + (rem == 0 ? 0.0 : rem <= sms ? 0.5 : 1.0);
    return small < big;
}

template <bool AK, bool BKM, int MI>
int launch_dgemm_mi(const GemmArgs &g)
{
    constexpr int BM = Cfg<MI>::BM, SMEM_BYTES = Cfg<MI>::SMEM_BYTES;
    CUtensorMap mapA, mapB;
    // K-major operand: box = 16 (k) x 128 rows; MN-major operand: box = 16 (m|n) x 16 (k)
    if (!tma::make_2d(&mapA, CU_TENSOR_MAP_DATA_TYPE_FLOAT64, 8, g.A, (uint64_t)g.rowsA, (uint64_t)g.colsA,
                      AK ? BM : BK, 16, (uint64_t)g.lda * 8) ||
        !tma::make_2d(&mapB, CU_TENSOR_MAP_DATA_TYPE_FLOAT64, 8, g.B, (uint64_t)g.rowsB, (uint64_t)g.colsB,
                      BKM ? BN : BK, 16, (uint64_t)g.ldb * 8))
        return cuda_fail(cudaErrorInvalidValue, "cuTensorMapEncodeTiled");
    static thread_local bool attr_done = false;
    if (!attr_done) {
        ACC_CUDA(cudaFuncSetAttribute(dgemm_dmma_kernel<AK, BKM, MI>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_BYTES));
        attr_done = true;
    }
    DgemmParams p;
    p.m = g.m; p.n = g.n; p.k = g.k; p.ldc = g.ldc;
    p.alpha = *static_cast<const double *>(g.alpha);
    p.C = static_cast<double *>(g.C);
    p.tiles_m = (int)((g.m + BM - 1) / BM);
    p.tiles_n = (int)((g.n + BN - 1) / BN);
    const int64_t tiles = (int64_t)p.tiles_m * p.tiles_n;
    if (tiles > 0x7fffffff) return ACCBLAS_ERR_SHAPE;
    dgemm_dmma_kernel<AK, BKM, MI><<<(unsigned)tiles, THREADS, SMEM_BYTES, g.stream>>>(mapA, mapB, p);
    return check_launch("dgemm_dmma_kernel");
}

template <bool AK, bool BKM>
int launch_dgemm(const GemmArgs &g)
{
    return prefer_small_tile(g.m, g.n) ? launch_dgemm_mi<AK, BKM, 4>(g) : launch_dgemm_mi<AK, BKM, 8>(g);
}

template <bool AK, bool BKM, bool CA, bool CB>
int launch_zgemm(const GemmArgs &g)
{
    CUtensorMap mapA, mapB;
    // maps are over doubles (no complex TMA type): a complex row of c columns is 2c doubles
    if (!tma::make_2d(&mapA, CU_TENSOR_MAP_DATA_TYPE_FLOAT64, 8, g.A, (uint64_t)g.rowsA, 2 * (uint64_t)g.colsA,
                      AK ? ZBM : ZBK, 16, (uint64_t)g.lda * 16) ||
        !tma::make_2d(&mapB, CU_TENSOR_MAP_DATA_TYPE_FLOAT64, 8, g.B, (uint64_t)g.rowsB, 2 * (uint64_t)g.colsB,
                      BKM ? ZBN : ZBK, 16, (uint64_t)g.ldb * 16))
        return cuda_fail(cudaErrorInvalidValue, "cuTensorMapEncodeTiled");
    static thread_local bool attr_done = false;
    if (!attr_done) {
        ACC_CUDA(cudaFuncSetAttribute(zgemm_dmma_kernel<AK, BKM, CA, CB>, cudaFuncAttributeMaxDynamicSharedMemorySize, ZSMEM_BYTES));
        attr_done = true;
    }
    ZgemmParams p;
    p.m = g.m; p.n = g.n; p.k = g.k; p.ldc = g.ldc;
    p.alr = static_cast<const double *>(g.alpha)[0];
    p.ali = static_cast<const double *>(g.alpha)[1];
    p.C = static_cast<double2 *>(g.C);
    p.tiles_m = (int)((g.m + ZBM - 1) / ZBM);
    p.tiles_n = (int)((g.n + ZBN - 1) / ZBN);
    const int64_t tiles = (int64_t)p.tiles_m * p.tiles_n;
    if (tiles > 0x7fffffff) return ACCBLAS_ERR_SHAPE;
    zgemm_dmma_kernel<AK, BKM, CA, CB><<<(unsigned)tiles, ZTHREADS, ZSMEM_BYTES, g.stream>>>(mapA, mapB, p);
    return check_launch("zgemm_dmma_kernel");
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
