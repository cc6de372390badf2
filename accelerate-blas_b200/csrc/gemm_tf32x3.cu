// gemm_tf32x3.cu -- FP32 / complex-FP32 GEMM on the 5th-generation tensor cores (tcgen05.mma
// kind::tf32, accumulators in TMEM) with FP32-class accuracy through a split-precision scheme.
// Replaces BLAS.sgemm / BLAS.cgemm at LLVM/PTX/Level3.hs:91,101.  The reference's cuBLAS call runs
// in the default math mode (true FP32), so plain TF32 would not be parity.
//
//   a = a_hi + a_lo,  a_hi = rna_tf32(a),  a_lo = rna_tf32(a - a_hi)      (|a - a_hi - a_lo| <= 2^-22 |a|)
//   a*b ~= a_hi*b_hi + a_hi*b_lo + a_lo*b_hi                               (dropped a_lo*b_lo <= 2^-22 |ab|)
//
// three tensor-core products per K-step, all accumulated in one FP32 TMEM accumulator.
//
// Two kernels:
//  1. pack_split: one O(n^2) pass per operand that (a) applies the op -- transposes to K-major,
//     conjugates -- (b) for complex types writes the REAL embedding
//         C_real (m x 2n) = A_real (m x 2k) * Bhat (2k x 2n),
//         Bhat[2l][2j] = br, Bhat[2l+1][2j] = -bi, Bhat[2l][2j+1] = bi, Bhat[2l+1][2j+1] = br
//     ("4M": 8mnk real flops, result lands directly in interleaved (re,im) layout, no epilogue
//     pass), and (c) splits every value into its hi / lo TF32 planes.
//     Output: P_hi, P_lo, each R x Kp row-major with pitch Kp rounded up to 32 floats (K-major for
//     both operands, so the GEMM kernel has exactly one layout).
//  2. sgemm_tf32x3_kernel: TMA (128-byte swizzle) -> 3-stage smem ring -> tcgen05.mma issued by one
//     thread -> TMEM accumulator -> tcgen05.ld epilogue, warp-specialised
//     (warp 0 TMA producer, warp 1 MMA issuer, warp 2 TMEM allocator, warps 4-7 epilogue).
//
// Workspace (stream-ordered, cudaMallocAsync on the caller's stream): 2 * 4 * (R_A + R_B) * Kp bytes.
#include "common.cuh"
#include "refit.cuh"
#include "tma.cuh"
#include <cstdlib>
#include <algorithm>

namespace accblas {
namespace {

// ------------------------------------------------------------------------------------
// pack_split
// ------------------------------------------------------------------------------------
enum PackMode {
    PK_REAL = 0,    // P[r][c] = X_op[r][c]
    PK_CPLX_A = 1,  // P[i][2l+p] = part p of A_op[i][l]          (imag * sign for H)
    PK_CPLX_B = 2   // P[2j+q][2l+p] = Bhat^T                       (see header)
};

struct PackArgs {
    const float *src;      // stored matrix (float view: complex = 2 floats per element)
    int64_t rows, cols;    // stored shape in ELEMENTS
    int64_t ld;            // stored row pitch in ELEMENTS
    int transposed;        // 1: op-index (r,l) lives at stored [l][r];  0: at stored [r][l]
    float conj_sign;       // +1, or -1 when the op is H on a complex type
    int64_t R, Kp;         // output extent (floats)
    int64_t pitch;         // output row pitch (floats), multiple of 32
    float *hi, *lo;
};

__device__ __forceinline__ float to_tf32(float x) {
    uint32_t r;
    asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(r) : "f"(x));
    return __uint_as_float(r);
}

template <int MODE>
__device__ __forceinline__ float pack_value(const PackArgs &a, int64_t r, int64_t c)
{
    if (MODE == PK_REAL) {
        return a.transposed ? a.src[c * a.ld + r] : a.src[r * a.ld + c];
    } else if (MODE == PK_CPLX_A) {
        const int64_t l = c >> 1;
        const int p = (int)(c & 1);
        const int64_t e = a.transposed ? l * a.ld + r : r * a.ld + l;
        const float v = a.src[2 * e + p];
        return p ? v * a.conj_sign : v;
    } else {
        const int64_t j = r >> 1, l = c >> 1;
        const int q = (int)(r & 1), p = (int)(c & 1);
        // op(B)[l][j]: stored [l][j] when op = N (not "transposed" w.r.t. k-major view means
        // stored [j][l]); `transposed` here means the K-major row j is NOT contiguous in memory
        const int64_t e = a.transposed ? l * a.ld + j : j * a.ld + l;
        const float br = a.src[2 * e], bi = a.src[2 * e + 1] * a.conj_sign;
        return q == p ? br : (q == 0 ? -bi : bi);
    }
}

// 32 x 32 output tile per 256-thread block.  When the source is contiguous along the output's
// row index (transposed), the tile goes through shared memory so both sides stay coalesced.
// Both operands are packed by ONE launch (blockIdx.z = 0: A, 1: B): the mid-size products this family
// now serves (1000^3 takes ~40 us in all) are sensitive to every extra launch.
// A block packs 32 rows x (PK_NT * 32) columns: all its loads are issued before the first store, so a
// thread keeps PK_NT * 4 loads in flight (the 32 x 32 version kept 4 and sat at 4.2 TB/s, latency-bound).
constexpr int PK_NT = 4;

template <int MODE>
__device__ __forceinline__ void pack_split_tile(const PackArgs &a, float (&tile)[PK_NT][32][33])
{
    const int64_t r0 = (int64_t)blockIdx.y * 32, c0 = (int64_t)blockIdx.x * (32 * PK_NT);
    if (r0 >= a.R || c0 >= a.pitch) return;                      // uniform per block (grid covers the larger operand)
    const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;    // 32 x 8
    float v[PK_NT][4];
    if (a.transposed) {
#pragma unroll
        for (int t = 0; t < PK_NT; ++t)
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                const int cc = ty + 8 * i;                       // fast index tx walks r
                const int64_t r = r0 + tx, c = c0 + 32 * t + cc;
                v[t][i] = (r < a.R && c < a.Kp) ? pack_value<MODE>(a, r, c) : 0.f;
            }
#pragma unroll
        for (int t = 0; t < PK_NT; ++t)
#pragma unroll
            for (int i = 0; i < 4; ++i) tile[t][ty + 8 * i][tx] = v[t][i];
        __syncthreads();
#pragma unroll
        for (int t = 0; t < PK_NT; ++t)
#pragma unroll
            for (int i = 0; i < 4; ++i) v[t][i] = tile[t][tx][ty + 8 * i];
    } else {
#pragma unroll
        for (int t = 0; t < PK_NT; ++t)
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                const int64_t r = r0 + ty + 8 * i, c = c0 + 32 * t + tx;   // fast index tx walks c
                v[t][i] = (r < a.R && c < a.Kp) ? pack_value<MODE>(a, r, c) : 0.f;
            }
    }
#pragma unroll
    for (int t = 0; t < PK_NT; ++t)
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            const int64_t r = r0 + ty + 8 * i, c = c0 + 32 * t + tx;
            if (r >= a.R || c >= a.pitch) continue;
            const float h = to_tf32(v[t][i]);
            a.hi[r * a.pitch + c] = h;
            a.lo[r * a.pitch + c] = to_tf32(v[t][i] - h);
        }
}

template <int MODE_A, int MODE_B>
__global__ void __launch_bounds__(256)
pack_split_kernel(const PackArgs a, const PackArgs b)
{
    __shared__ float tile[PK_NT][32][33];
    if (blockIdx.z == 0) pack_split_tile<MODE_A>(a, tile);
    else pack_split_tile<MODE_B>(b, tile);
}

int launch_pack(bool cplx, const PackArgs &a, const PackArgs &b, cudaStream_t st)
{
    dim3 grid((unsigned)((std::max(a.pitch, b.pitch) + 32 * PK_NT - 1) / (32 * PK_NT)), (unsigned)((std::max(a.R, b.R) + 31) / 32), 2);
    if (grid.y > 65535u) return ACCBLAS_ERR_SHAPE;     // eligibility keeps R below 2^21
    if (cplx) pack_split_kernel<PK_CPLX_A, PK_CPLX_B><<<grid, 256, 0, st>>>(a, b);
    else      pack_split_kernel<PK_REAL, PK_REAL><<<grid, 256, 0, st>>>(a, b);
    return check_launch("pack_split_kernel");
}


// ------------------------------------------------------------------------------------
// 3M for complex-float products: op(X) = Xr + i Xi is packed as THREE real K-major operands
//   P0 = Xr,  P1 = s * Xi,  P2 = Xr + s * Xi      (s = -1 under H), each split into hi / lo planes,
// so that C = alpha * ((P0a P0b - P1a P1b) + i (P2a P2b - P0a P0b - P1a P1b)) takes three real products
// of extent m x n x k instead of the one m x 2n x 2k product of the real embedding (6 mnk against 8 mnk).
// Same 32 x 32 tiling as pack_split; R / Kp / pitch count complex elements here.
// ------------------------------------------------------------------------------------
struct Pack3Args {
    const float2 *src;     // stored complex matrix
    int64_t ld;            // stored row pitch (complex elements)
    int transposed;        // 1: op-index (r,l) lives at stored [l][r]
    float conj_sign;
    int64_t R, Kp, pitch;  // output extent and row pitch (floats), pitch multiple of 32
    float *hi[3], *lo[3];
};

__device__ __forceinline__ void pack_split3_tile(const Pack3Args &a, float2 (&tile)[32][33])
{
    const int64_t r0 = (int64_t)blockIdx.y * 32, c0 = (int64_t)blockIdx.x * 32;
    if (r0 >= a.R || c0 >= a.pitch) return;
    const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
    if (a.transposed) {
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            const int cc = ty + 8 * i;
            const int64_t r = r0 + tx, c = c0 + cc;
            tile[cc][tx] = (r < a.R && c < a.Kp) ? a.src[c * a.ld + r] : make_float2(0.f, 0.f);
        }
        __syncthreads();
    }
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const int rr = ty + 8 * i;
        const int64_t r = r0 + rr, c = c0 + tx;
        if (r >= a.R || c >= a.pitch) continue;
        float2 z = make_float2(0.f, 0.f);
        if (c < a.Kp) z = a.transposed ? tile[tx][rr] : a.src[r * a.ld + c];
        const float im = z.y * a.conj_sign;
        const float v[3] = {z.x, im, z.x + im};
#pragma unroll
        for (int p = 0; p < 3; ++p) {
            const float h = to_tf32(v[p]);
            a.hi[p][r * a.pitch + c] = h;
            a.lo[p][r * a.pitch + c] = to_tf32(v[p] - h);
        }
    }
}

__global__ void __launch_bounds__(256) pack_split3_kernel(const Pack3Args a, const Pack3Args b)
{
    __shared__ float2 tile[32][33];
    pack_split3_tile(blockIdx.z == 0 ? a : b, tile);
}

struct Combine3Args {
    const float *t[3];     // m x n real products, row pitch ldt (multiple of 4 floats)
    float2 *C;
    int64_t m, n, ldt, ldc;
    Alpha<float2> alpha;
    BetaC<float2> beta;
    Refit32 refit;         // non-finite results are recomputed from the caller's operands (refit.cuh)
};

// one thread = four consecutive columns: three 16-byte loads, two 16-byte stores (when C allows)
__global__ void __launch_bounds__(256) combine3_kernel(const Combine3Args a)
{
    const float2 al = a.alpha.get();
    const int64_t quads = a.ldt / 4, total = a.m * quads;
    const bool vec = (a.ldc & 1) == 0 && (reinterpret_cast<uintptr_t>(a.C) & 15u) == 0;
    // the loop bound is uniform per warp (the refit votes need converged warps): `live` marks the threads with work
    for (int64_t base = (int64_t)blockIdx.x * blockDim.x; base < total; base += (int64_t)gridDim.x * blockDim.x) {
        const int64_t i = base + threadIdx.x;
        const int64_t r = i < total ? i / quads : 0, c = i < total ? (i - r * quads) * 4 : 0;
        const bool live = i < total && c < a.n;
        const int64_t o = r * a.ldt + c;
        float4 t1 = make_float4(0.f, 0.f, 0.f, 0.f), t2 = t1, t3 = t1;
        if (live) {
            t1 = *reinterpret_cast<const float4 *>(a.t[0] + o);
            t2 = *reinterpret_cast<const float4 *>(a.t[1] + o);
            t3 = *reinterpret_cast<const float4 *>(a.t[2] + o);
        }
        const float x1[4] = {t1.x, t1.y, t1.z, t1.w}, x2[4] = {t2.x, t2.y, t2.z, t2.w}, x3[4] = {t3.x, t3.y, t3.z, t3.w};
        float2 out[4];
        bool any = false;
#pragma unroll
        for (int e = 0; e < 4; ++e) {
            out[e] = make_float2(x1[e] - x2[e], (x3[e] - x1[e]) - x2[e]);
            any |= nonfinite32(out[e].x) || nonfinite32(out[e].y);
        }
        if (__any_sync(0xffffffffu, any && live)) {
#pragma unroll
            for (int e = 0; e < 4; ++e) out[e] = refit_cplx(a.refit, out[e], r, c + e, live && c + e < a.n);
        }
        if (!live) continue;
#pragma unroll
        for (int e = 0; e < 4; ++e) {
            const float xr = out[e].x, xi = out[e].y;
            out[e] = make_float2(al.x * xr - al.y * xi, al.x * xi + al.y * xr);
            if (a.beta.on() && c + e < a.n) out[e] = a.beta.add(out[e], r, c + e);
        }
        float2 *dst = a.C + r * a.ldc + c;
        if (vec && c + 4 <= a.n) {
            *reinterpret_cast<float4 *>(dst) = make_float4(out[0].x, out[0].y, out[1].x, out[1].y);
            *reinterpret_cast<float4 *>(dst + 2) = make_float4(out[2].x, out[2].y, out[3].x, out[3].y);
        } else {
#pragma unroll
            for (int e = 0; e < 4; ++e)
                if (c + e < a.n) dst[e] = out[e];
        }
    }
}

// ------------------------------------------------------------------------------------
// tcgen05 helpers (inline PTX; SASS: UTCHMMA / LDTM / UTCBAR)
// ------------------------------------------------------------------------------------
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tmem_alloc(uint32_t slot_smem, uint32_t ncols) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(slot_smem), "r"(ncols) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc(uint32_t taddr, uint32_t ncols) {
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(ncols) : "memory");
}
// D[tmem] (+)= A[smem desc] * B[smem desc], kind::tf32, issued by ONE thread
__device__ __forceinline__ void umma_tf32(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n\t}"
        ::"r"(tmem_d), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate) : "memory");
}
// arrive on an mbarrier when all MMAs issued so far by this thread have completed
__device__ __forceinline__ void umma_commit(uint32_t bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void tmem_ld32(uint32_t taddr, uint32_t (&v)[32]) {
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
        "{%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,"
        "%16,%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31}, [%32];"
        : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]),
          "=r"(v[8]), "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15]),
          "=r"(v[16]), "=r"(v[17]), "=r"(v[18]), "=r"(v[19]), "=r"(v[20]), "=r"(v[21]), "=r"(v[22]), "=r"(v[23]),
          "=r"(v[24]), "=r"(v[25]), "=r"(v[26]), "=r"(v[27]), "=r"(v[28]), "=r"(v[29]), "=r"(v[30]), "=r"(v[31])
        : "r"(taddr) : "memory");
}
__device__ __forceinline__ void tmem_ld_wait() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }

// K-major, 128-byte-swizzled operand tile: rows of 128 bytes, 8-row groups 1024 bytes apart.
// (cute::UMMA::SmemDescriptor: start>>4 [0,14), LBO>>4 [16,30) = 1, SBO>>4 [32,46) = 64,
//  version [46,48) = 1, layout_type [61,64) = 2 = SWIZZLE_128B)
__device__ __forceinline__ uint64_t smem_desc_sw128(uint32_t saddr) {
    return (uint64_t)((saddr >> 4) & 0x3fffu) | (1ull << 16) | (64ull << 32) | (1ull << 46) | (2ull << 61);
}
// cute::UMMA::InstrDescriptor: c_format F32 (1) [4,6), a/b_format TF32 (2) [7,10) [10,13),
// a/b K-major (0) [15],[16], N>>3 [17,23), M>>4 [24,29)
__host__ __device__ constexpr uint32_t idesc_tf32(int M, int N) {
    return (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(N >> 3) << 17) | ((uint32_t)(M >> 4) << 24);
}

// ------------------------------------------------------------------------------------
// sgemm_tf32x3_kernel: C_real (M x N) = alpha * P_A (M x Kp) * P_B^T (N x Kp), both K-major
//
// Accuracy note.  The tensor core adds each K = 8 product group into the FP32 TMEM accumulator
// with truncation, so a long accumulation chain drifts (measured: ~k/8 * 3 * ulp(acc)/2, i.e.
// 1.6e-3 absolute at k = 2048 -- far inside the normwise tolerance but ~50x a true FP32 sum).
// The chain is therefore cut every KCB K-blocks ("promotion"): the MMA warp alternates between
// two TMEM accumulators, starting each chunk from zero, and the eight epilogue warps drain the
// finished chunk with tcgen05.ld and add it into FP32 REGISTER accumulators with round-to-nearest
// FADDs while the next chunk is being multiplied.  Error then scales like a blocked FP32 sum.
// ------------------------------------------------------------------------------------
constexpr int TBM = 128, TBK = 32;          // 32 floats = one 128-byte swizzle row
constexpr int TTHREADS = 384;               // warps 0-3 control warpgroup, warps 4-11 epilogue
constexpr int EPI_WARPS = 8;

struct SgemmParams {
    int64_t M, N;          // real extents of C_real
    int64_t Kp;            // packed inner extent
    int64_t ldc;           // floats
    float *C;
    Alpha<float2> alpha;   // (re, im); im = 0 for real
    BetaC<float> beta;     // + beta * Cin on the float view (real types; complex beta: cbeta over the same floats)
    float2 cbeta;          // complex beta (cplx = 1), applied to (re, im) column pairs of Cin
    int cplx;              // 1: columns are (re,im) pairs, alpha is complex
    int tiles_m, tiles_n;
    // plane batching (3M): the grid holds `planes` independent products of the same extents whose operands are
    // stacked along the row dimension of the tensor maps; 0 tiles_per_plane = a single product
    int tiles_per_plane;
    int plane_rows_a, plane_rows_b;      // row offset of plane p in the A / B maps: p * plane_rows_*
    int64_t plane_stride_c;              // floats between the C planes
    Refit32 refit;                       // non-finite accumulators are recomputed in plain FP32 (refit.cuh); A == nullptr: off
};

template <int TBN, int STAGES>
struct TcCfg {
    static constexpr int A_TILE = TBM * TBK * 4;               // 16 KiB
    static constexpr int B_TILE = TBN * TBK * 4;
    static constexpr int STAGE = 2 * A_TILE + 2 * B_TILE;      // hi + lo of both operands
    static constexpr int BAR_OFF = STAGES * STAGE;
    static constexpr int SMEM = STAGES * STAGE + 1024 + 256;
};

// ---- C = alpha * acc (+ beta * Cin): one thread owns HALF consecutive columns of one row ----
// (called by whole warps: the non-finite refit votes)
template <int HALF>
__device__ __forceinline__ void tc_store(const SgemmParams &p, float (&acc)[HALF], int64_t row, int64_t col0, int plane)
{
    if (p.refit.A) {                                           // non-finite sums: see refit.cuh (warp-uniform)
        bool any = false;
#pragma unroll
        for (int i = 0; i < HALF; ++i) any |= nonfinite32(acc[i]);
        if (__any_sync(0xffffffffu, any && row < p.M)) {
            if (p.cplx) {
#pragma unroll
                for (int i = 0; i < HALF; i += 2) {
                    const float2 v = refit_cplx(p.refit, make_float2(acc[i], acc[i + 1]), row, (col0 + i) >> 1,
                                                row < p.M && col0 + i < p.N);
                    acc[i] = v.x; acc[i + 1] = v.y;
                }
            } else {
#pragma unroll
                for (int i = 0; i < HALF; ++i)
                    acc[i] = refit_real(p.refit, acc[i], row, col0 + i, row < p.M && col0 + i < p.N);
            }
        }
    }
    const float2 al = p.alpha.dev ? (p.cplx ? *p.alpha.dev : make_float2(*reinterpret_cast<const float *>(p.alpha.dev), 0.f))
                                  : p.alpha.v;
    if (row < p.M) {
        float *crow = p.C + plane * p.plane_stride_c + row * p.ldc;
        const bool vec4 = (p.ldc & 3) == 0 && ((reinterpret_cast<uintptr_t>(p.C) & 15u) == 0) && col0 + HALF <= p.N;
        if (p.cplx) {
#pragma unroll
            for (int i = 0; i < HALF; i += 2) {
                const float xr = acc[i], xi = acc[i + 1];
                acc[i] = al.x * xr - al.y * xi;
                acc[i + 1] = al.x * xi + al.y * xr;
            }
        } else {
#pragma unroll
            for (int i = 0; i < HALF; ++i) acc[i] *= al.x;
        }
        if (p.beta.on()) {                                     // + beta * Cin (accblas_gemm_ex): one row of Cin per thread
            const float *cin = p.beta.cin + row * p.beta.ld + col0;
            if (p.cplx) {
#pragma unroll
                for (int i = 0; i < HALF; i += 2)
                    if (col0 + i < p.N) {                      // N is even for complex: a started pair is whole
                        const float cr = cin[i], ci = cin[i + 1];
                        acc[i] += p.cbeta.x * cr - p.cbeta.y * ci;
                        acc[i + 1] += p.cbeta.x * ci + p.cbeta.y * cr;
                    }
            } else {
#pragma unroll
                for (int i = 0; i < HALF; ++i)
                    if (col0 + i < p.N) acc[i] = fmaf(p.beta.beta, cin[i], acc[i]);
            }
        }
        if (vec4) {
#pragma unroll
            for (int i = 0; i < HALF; i += 4)
                *reinterpret_cast<float4 *>(crow + col0 + i) = make_float4(acc[i], acc[i + 1], acc[i + 2], acc[i + 3]);
        } else {
#pragma unroll
            for (int i = 0; i < HALF; ++i)
                if (col0 + i < p.N) crow[col0 + i] = acc[i];
        }
    }
}

template <int TBN, int STAGES, int KCB>
__global__ void __launch_bounds__(TTHREADS, 1)
sgemm_tf32x3_kernel(const __grid_constant__ CUtensorMap mapAhi, const __grid_constant__ CUtensorMap mapAlo,
                    const __grid_constant__ CUtensorMap mapBhi, const __grid_constant__ CUtensorMap mapBlo,
                    const SgemmParams p)
{
    using Cfg = TcCfg<TBN, STAGES>;
    constexpr int HALF = TBN / 2;                                  // columns per epilogue warp
    extern __shared__ uint8_t smem_raw[];
    const uint32_t smem_base = (tma::smem_u32(smem_raw) + 1023u) & ~1023u;
    const uint32_t bar_full = smem_base + Cfg::BAR_OFF;            // full[STAGES]
    const uint32_t bar_empty = bar_full + 8 * STAGES;              // empty[STAGES]
    const uint32_t bar_accf = bar_empty + 8 * STAGES;              // acc_full[2]  (MMA -> epilogue)
    const uint32_t bar_acce = bar_accf + 16;                       // acc_empty[2] (epilogue -> MMA)
    const uint32_t tmem_slot = bar_acce + 16;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

    constexpr int GROUP = 8;
    int tile_m, tile_n, plane = 0, bid;
    {
        bid = blockIdx.x;
        if (p.tiles_per_plane > 0) {
            plane = bid / p.tiles_per_plane;
            bid -= plane * p.tiles_per_plane;
        }
        const int per_group = GROUP * p.tiles_n;
        const int grp = bid / per_group;
        const int first_m = grp * GROUP;
        const int gsz = min(p.tiles_m - first_m, GROUP);
        const int in_grp = bid - grp * per_group;
        tile_m = first_m + in_grp % gsz;
        tile_n = in_grp / gsz;
    }
    const int m0 = tile_m * TBM, n0 = tile_n * TBN;
    const int ma = m0 + plane * p.plane_rows_a, nb = n0 + plane * p.plane_rows_b;   // rows in the stacked maps
    const int kblocks = (int)((p.Kp + TBK - 1) / TBK);
    const int chunks = (kblocks + KCB - 1) / KCB;

    if (threadIdx.x == 0) {
        for (int s = 0; s < STAGES; ++s) {
            tma::mbar_init(bar_full + 8 * s, 1);
            tma::mbar_init(bar_empty + 8 * s, 1);
        }
        for (int b = 0; b < 2; ++b) {
            tma::mbar_init(bar_accf + 8 * b, 1);
            tma::mbar_init(bar_acce + 8 * b, EPI_WARPS);
        }
        tma::fence_barrier_init();
    }
    if (warp == 2) tmem_alloc(tmem_slot, 2 * TBN);                 // two accumulators of TBN fp32 columns
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    uint32_t tmem_base;
    asm volatile("ld.shared.u32 %0, [%1];" : "=r"(tmem_base) : "r"(tmem_slot));

    if (warp < 4) {
        asm volatile("setmaxnreg.dec.sync.aligned.u32 40;");
        if (warp == 0 && lane == 0) {
            // ================================ TMA producer ================================
            tma::prefetch_tmap(&mapAhi); tma::prefetch_tmap(&mapAlo);
            tma::prefetch_tmap(&mapBhi); tma::prefetch_tmap(&mapBlo);
            for (int kb = 0; kb < kblocks; ++kb) {
                const int s = kb % STAGES;
                if (kb >= STAGES) tma::mbar_wait(bar_empty + 8 * s, ((kb / STAGES) - 1) & 1);
                const uint32_t full = bar_full + 8 * s;
                const uint32_t st = smem_base + s * Cfg::STAGE;
                tma::mbar_expect_tx(full, Cfg::STAGE);
                const int k0 = kb * TBK;
                tma::load_2d(st, &mapAhi, full, k0, ma);
                tma::load_2d(st + Cfg::A_TILE, &mapAlo, full, k0, ma);
                tma::load_2d(st + 2 * Cfg::A_TILE, &mapBhi, full, k0, nb);
                tma::load_2d(st + 2 * Cfg::A_TILE + Cfg::B_TILE, &mapBlo, full, k0, nb);
            }
        } else if (warp == 1) {
            // ================================ MMA issuer ================================
            constexpr uint32_t idesc = idesc_tf32(TBM, TBN);
            int kb = 0;
            for (int c = 0; c < chunks; ++c) {
                const int b = c & 1;
                if (c >= 2) {                                       // accumulator b drained by the epilogue?
                    tma::mbar_wait(bar_acce + 8 * b, ((c >> 1) - 1) & 1);
                    tc_fence_after();
                }
                const uint32_t tacc = tmem_base + (uint32_t)(b * TBN);
                const int kend = min(kb + KCB, kblocks);
                for (int first = 1; kb < kend; ++kb) {
                    const int s = kb % STAGES;
                    tma::mbar_wait(bar_full + 8 * s, (kb / STAGES) & 1);
                    tc_fence_after();
                    if (lane == 0) {
                        const uint32_t st = smem_base + s * Cfg::STAGE;
                        const uint64_t ahi = smem_desc_sw128(st), alo = smem_desc_sw128(st + Cfg::A_TILE);
                        const uint64_t bhi = smem_desc_sw128(st + 2 * Cfg::A_TILE);
                        const uint64_t blo = smem_desc_sw128(st + 2 * Cfg::A_TILE + Cfg::B_TILE);
#pragma unroll
                        for (int k = 0; k < TBK / 8; ++k) {       // UMMA_K = 8 tf32 = 32 bytes = +2 in the address field
                            const uint64_t adv = (uint64_t)(2 * k);
                            umma_tf32(tacc, alo + adv, bhi + adv, idesc, (first && k == 0) ? 0u : 1u);
                            umma_tf32(tacc, ahi + adv, blo + adv, idesc, 1);
                            umma_tf32(tacc, ahi + adv, bhi + adv, idesc, 1);
                        }
                        umma_commit(bar_empty + 8 * s);            // stage reusable once these MMAs retire
                    }
                    first = 0;
                    __syncwarp();
                }
                if (lane == 0) umma_commit(bar_accf + 8 * b);      // chunk accumulator complete
                __syncwarp();
            }
        }
    } else {
        // ================================ promotion + epilogue ================================
        asm volatile("setmaxnreg.inc.sync.aligned.u32 232;");
        const int e = warp - 4;
        const int q = e & 3;                                       // TMEM lane quadrant (warp id % 4)
        const int h = e >> 2;                                      // column half
        float acc[HALF];
#pragma unroll
        for (int i = 0; i < HALF; ++i) acc[i] = 0.f;
        for (int c = 0; c < chunks; ++c) {
            const int b = c & 1;
            tma::mbar_wait(bar_accf + 8 * b, (c >> 1) & 1);
            tc_fence_after();
            const uint32_t t0 = tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(b * TBN + h * HALF);
            if constexpr (HALF >= 64) {
#pragma unroll
                for (int c0 = 0; c0 < HALF; c0 += 64) {
                    uint32_t v0[32], v1[32];
                    tmem_ld32(t0 + c0, v0);
                    tmem_ld32(t0 + c0 + 32, v1);
                    tmem_ld_wait();
#pragma unroll
                    for (int i = 0; i < 32; ++i) {
                        acc[c0 + i] += __uint_as_float(v0[i]);
                        acc[c0 + 32 + i] += __uint_as_float(v1[i]);
                    }
                }
            } else {                                               // 128 x 64 tile: 32 columns per warp
                uint32_t v0[32];
                tmem_ld32(t0, v0);
                tmem_ld_wait();
#pragma unroll
                for (int i = 0; i < 32; ++i) acc[i] += __uint_as_float(v0[i]);
            }
            tc_fence_before();
            __syncwarp();
            if (lane == 0) tma::mbar_arrive(bar_acce + 8 * b);
        }
        tc_store<HALF>(p, acc, m0 + q * 32 + lane, n0 + h * HALF, plane);
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 2) {
        tc_fence_after();
        tmem_dealloc(tmem_base, 2 * TBN);
    }
}

// ------------------------------------------------------------------------------------
// sgemm_tf32x3_pair_kernel: the same product on CTA PAIRS (tcgen05.mma.cta_group::2, UMMA M = 256).
//
// Why: per K-block the one-CTA 128 x 256 tile moves (128 + 256) x 32 floats x {hi, lo} = 96 KiB through the SM's
// shared memory once as TMA writes and 1.5 times as tensor-core operand reads (three MMAs per K-step) -- 240 KiB per
// 1536 tensor clocks = 156 B/clk against the 128 B/clk the SM's shared memory delivers: the pipe cannot be more than
// ~82 % busy (measured round 1: 80 %).  A pair of CTAs on the two SMs of a TPC computes a 256 x 256 tile: each CTA
// stages its own 128 rows of A and only HALF of B (128 of the 256 columns), the hardware feeds both tensor cores
// with both halves, so per CTA a K-block is 64 KiB of TMA writes + 96 KiB of operand reads = 104 B/clk, and the
// L2 -> SM traffic per flop drops by a third as well.  Three stages of 64 KiB replace two of 96 KiB.
//
// Protocol (rank = %cluster_ctarank; rank 0 = leader):
//   full[s]   lives in the LEADER: one arrive.expect_tx by the leader's producer for the bytes of BOTH CTAs; each CTA's
//             producer issues its own TMA loads (.cta_group::2) that complete_tx on the leader's barrier;
//   empty[s]  one per CTA, released by the leader's tcgen05.commit multicast to both CTAs;
//   acc_full  one per CTA (commit multicast); acc_empty in the leader, 2 x 8 epilogue-warp arrivals (the peer's are
//             remote arrives with cluster-scope release);
//   TMEM      allocated / freed by warp 2 of BOTH CTAs (cta_group::2), cluster barrier before and after.
// The promotion scheme (two accumulators, chunks of KCB K-blocks drained into FP32 registers) is unchanged.
// ------------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t cluster_ctarank() {
    uint32_t r;
    asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
    return r;
}
__device__ __forceinline__ void cluster_sync_all() {
    asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
    asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
}
__device__ __forceinline__ uint32_t map_to_rank(uint32_t saddr, uint32_t rank) {
    uint32_t r;
    asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(saddr), "r"(rank));
    return r;
}
__device__ __forceinline__ void mbar_arrive_remote(uint32_t cluster_addr) {
    asm volatile("mbarrier.arrive.release.cluster.shared::cluster.b64 _, [%0];" ::"r"(cluster_addr) : "memory");
}
__device__ __forceinline__ void tmem_alloc2(uint32_t slot_smem, uint32_t ncols) {
    asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(slot_smem), "r"(ncols) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc2(uint32_t taddr, uint32_t ncols) {
    asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(ncols) : "memory");
}
__device__ __forceinline__ void umma_tf32_pair(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::2.kind::tf32 [%0], %1, %2, %3, p;\n\t}"
        ::"r"(tmem_d), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate) : "memory");
}
// arrive on the barrier at this shared-memory offset in BOTH CTAs of the pair once all MMAs issued so far retire
__device__ __forceinline__ void umma_commit_pair(uint32_t bar) {
    asm volatile("tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;"
                 ::"r"(bar), "h"((uint16_t)3) : "memory");
}
// 2-D tiled load into THIS CTA's shared memory, completing on a barrier of the pair's leader
__device__ __forceinline__ void tma_load_2d_pair(uint32_t dst, const void *tmap, uint32_t leader_bar, int c0, int c1) {
    asm volatile(
        "cp.async.bulk.tensor.2d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
        ::"r"(dst), "l"(tmap), "r"(leader_bar), "r"(c0), "r"(c1) : "memory");
}

constexpr int PBN = 256;                     // pair tile: 256 x 256, each CTA 128 rows x 256 columns of C
template <int STAGES>
struct PairCfg {
    static constexpr int A_TILE = TBM * TBK * 4;               // 16 KiB: this CTA's 128 rows of A
    static constexpr int B_TILE = (PBN / 2) * TBK * 4;         // 16 KiB: this CTA's half of the 256 columns of B
    static constexpr int STAGE = 2 * A_TILE + 2 * B_TILE;      // 64 KiB
    static constexpr int BAR_OFF = STAGES * STAGE;
    static constexpr int SMEM = STAGES * STAGE + 1024 + 256;
};

template <int STAGES, int KCB>
__global__ void __launch_bounds__(TTHREADS, 1)
sgemm_tf32x3_pair_kernel(const __grid_constant__ CUtensorMap mapAhi, const __grid_constant__ CUtensorMap mapAlo,
                         const __grid_constant__ CUtensorMap mapBhi, const __grid_constant__ CUtensorMap mapBlo,
                         const SgemmParams p)
{
    using Cfg = PairCfg<STAGES>;
    constexpr int HALF = PBN / 2;                                  // columns per epilogue warp
    extern __shared__ uint8_t smem_raw[];
    const uint32_t smem_base = (tma::smem_u32(smem_raw) + 1023u) & ~1023u;
    const uint32_t bar_full = smem_base + Cfg::BAR_OFF;            // full[STAGES]   (used in the leader only)
    const uint32_t bar_empty = bar_full + 8 * STAGES;              // empty[STAGES]  (one set per CTA)
    const uint32_t bar_accf = bar_empty + 8 * STAGES;              // acc_full[2]    (one set per CTA)
    const uint32_t bar_acce = bar_accf + 16;                       // acc_empty[2]   (used in the leader only)
    const uint32_t tmem_slot = bar_acce + 16;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const uint32_t rank = cluster_ctarank();
    const bool leader = rank == 0;

    constexpr int GROUP = 8;
    int tile_m, tile_n, plane = 0;
    {
        int bid = blockIdx.x >> 1;                                 // pair index (cluster = 2 consecutive CTAs)
        if (p.tiles_per_plane > 0) {
            plane = bid / p.tiles_per_plane;
            bid -= plane * p.tiles_per_plane;
        }
        const int per_group = GROUP * p.tiles_n;
        const int grp = bid / per_group;
        const int first_m = grp * GROUP;
        const int gsz = min(p.tiles_m - first_m, GROUP);
        const int in_grp = bid - grp * per_group;
        tile_m = first_m + in_grp % gsz;
        tile_n = in_grp / gsz;
    }
    const int m0 = tile_m * (2 * TBM) + (int)rank * TBM, n0 = tile_n * PBN;
    const int ma = m0 + plane * p.plane_rows_a;                    // this CTA's 128 rows of A in the stacked map
    const int nb = n0 + (int)rank * (PBN / 2) + plane * p.plane_rows_b;   // this CTA's half of the B rows
    const int kblocks = (int)((p.Kp + TBK - 1) / TBK);
    const int chunks = (kblocks + KCB - 1) / KCB;

    if (threadIdx.x == 0) {
        for (int s = 0; s < STAGES; ++s) {
            tma::mbar_init(bar_full + 8 * s, 1);
            tma::mbar_init(bar_empty + 8 * s, 1);
        }
        for (int b = 0; b < 2; ++b) {
            tma::mbar_init(bar_accf + 8 * b, 1);
            tma::mbar_init(bar_acce + 8 * b, 2 * EPI_WARPS);
        }
        tma::fence_barrier_init();
    }
    if (warp == 2) tmem_alloc2(tmem_slot, 2 * PBN);                // both CTAs, same warp: two accumulators of 256 columns
    tc_fence_before();
    cluster_sync_all();                                            // the peer's barriers exist before anything signals them
    tc_fence_after();
    uint32_t tmem_base;
    asm volatile("ld.shared.u32 %0, [%1];" : "=r"(tmem_base) : "r"(tmem_slot));

    if (warp < 4) {
        asm volatile("setmaxnreg.dec.sync.aligned.u32 40;");
        if (warp == 0 && lane == 0) {
            // ================================ TMA producer (both CTAs) ================================
            tma::prefetch_tmap(&mapAhi); tma::prefetch_tmap(&mapAlo);
            tma::prefetch_tmap(&mapBhi); tma::prefetch_tmap(&mapBlo);
            const uint32_t leader_full = map_to_rank(bar_full, 0);
            for (int kb = 0; kb < kblocks; ++kb) {
                const int s = kb % STAGES;
                if (kb >= STAGES) tma::mbar_wait(bar_empty + 8 * s, ((kb / STAGES) - 1) & 1);
                if (leader) tma::mbar_expect_tx(bar_full + 8 * s, 2 * Cfg::STAGE);     // both CTAs' bytes
                const uint32_t full = leader_full + 8 * s;
                const uint32_t st = smem_base + s * Cfg::STAGE;
                const int k0 = kb * TBK;
                tma_load_2d_pair(st, &mapAhi, full, k0, ma);
                tma_load_2d_pair(st + Cfg::A_TILE, &mapAlo, full, k0, ma);
                tma_load_2d_pair(st + 2 * Cfg::A_TILE, &mapBhi, full, k0, nb);
                tma_load_2d_pair(st + 2 * Cfg::A_TILE + Cfg::B_TILE, &mapBlo, full, k0, nb);
            }
        } else if (warp == 1 && leader) {
            // ================================ MMA issuer (leader only) ================================
            constexpr uint32_t idesc = idesc_tf32(2 * TBM, PBN);
            int kb = 0;
            for (int c = 0; c < chunks; ++c) {
                const int b = c & 1;
                if (c >= 2) {                                       // accumulator b drained by BOTH CTAs' epilogues?
                    tma::mbar_wait(bar_acce + 8 * b, ((c >> 1) - 1) & 1);
                    tc_fence_after();
                }
                const uint32_t tacc = tmem_base + (uint32_t)(b * PBN);
                const int kend = min(kb + KCB, kblocks);
                for (int first = 1; kb < kend; ++kb) {
                    const int s = kb % STAGES;
                    tma::mbar_wait(bar_full + 8 * s, (kb / STAGES) & 1);
                    tc_fence_after();
                    if (lane == 0) {
                        const uint32_t st = smem_base + s * Cfg::STAGE;
                        const uint64_t ahi = smem_desc_sw128(st), alo = smem_desc_sw128(st + Cfg::A_TILE);
                        const uint64_t bhi = smem_desc_sw128(st + 2 * Cfg::A_TILE);
                        const uint64_t blo = smem_desc_sw128(st + 2 * Cfg::A_TILE + Cfg::B_TILE);
#pragma unroll
                        for (int k = 0; k < TBK / 8; ++k) {
                            const uint64_t adv = (uint64_t)(2 * k);
                            umma_tf32_pair(tacc, alo + adv, bhi + adv, idesc, (first && k == 0) ? 0u : 1u);
                            umma_tf32_pair(tacc, ahi + adv, blo + adv, idesc, 1);
                            umma_tf32_pair(tacc, ahi + adv, bhi + adv, idesc, 1);
                        }
                        umma_commit_pair(bar_empty + 8 * s);       // the stage is reusable in both CTAs once these retire
                    }
                    first = 0;
                    __syncwarp();
                }
                if (lane == 0) umma_commit_pair(bar_accf + 8 * b); // chunk accumulator complete (both CTAs' epilogues)
                __syncwarp();
            }
        }
    } else {
        // ================================ promotion + epilogue (both CTAs) ================================
        asm volatile("setmaxnreg.inc.sync.aligned.u32 232;");
        const int e = warp - 4;
        const int q = e & 3;                                       // TMEM lane quadrant (warp id % 4)
        const int h = e >> 2;                                      // column half
        const uint32_t leader_acce = map_to_rank(bar_acce, 0);
        float acc[HALF];
#pragma unroll
        for (int i = 0; i < HALF; ++i) acc[i] = 0.f;
        for (int c = 0; c < chunks; ++c) {
            const int b = c & 1;
            tma::mbar_wait(bar_accf + 8 * b, (c >> 1) & 1);
            tc_fence_after();
            const uint32_t t0 = tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(b * PBN + h * HALF);
#pragma unroll
            for (int c0 = 0; c0 < HALF; c0 += 64) {
                uint32_t v0[32], v1[32];
                tmem_ld32(t0 + c0, v0);
                tmem_ld32(t0 + c0 + 32, v1);
                tmem_ld_wait();
#pragma unroll
                for (int i = 0; i < 32; ++i) {
                    acc[c0 + i] += __uint_as_float(v0[i]);
                    acc[c0 + 32 + i] += __uint_as_float(v1[i]);
                }
            }
            tc_fence_before();
            __syncwarp();
            if (lane == 0) {
                if (leader) tma::mbar_arrive(bar_acce + 8 * b);
                else mbar_arrive_remote(leader_acce + 8 * b);
            }
        }
        tc_store<HALF>(p, acc, (int64_t)m0 + q * 32 + lane, (int64_t)n0 + h * HALF, plane);
    }
    tc_fence_before();
    cluster_sync_all();                                            // the leader's MMAs have read the peer's shared memory
    if (warp == 2) {
        tc_fence_after();
        tmem_dealloc2(tmem_base, 2 * PBN);
    }
}

// ------------------------------------------------------------------------------------
// host side
// ------------------------------------------------------------------------------------
struct TcPlan {
    int64_t RA, RB, Kp, pitch;     // packed extents (floats)
    int64_t M, N;                  // C_real extents
    size_t plane_a, plane_b;       // bytes of one hi (or lo) plane
};

TcPlan make_plan(const GemmArgs &g)
{
    TcPlan t;
    const bool cx = g.dtype == ACCBLAS_C32;
    t.RA = g.m;
    t.RB = cx ? 2 * g.n : g.n;
    t.Kp = cx ? 2 * g.k : g.k;
    t.pitch = (t.Kp + 31) / 32 * 32;
    t.M = g.m;
    t.N = t.RB;
    t.plane_a = ((size_t)t.RA * (size_t)t.pitch * 4 + 1023) / 1024 * 1024;
    t.plane_b = ((size_t)t.RB * (size_t)t.pitch * 4 + 1023) / 1024 * 1024;
    return t;
}

// CTA tile choice (always 128 rows: UMMA M = 128 fills the 128 TMEM lanes).  A launch lasts
// ceil(tiles / SMs) waves of one tile per SM; the cost of a K = 8 step of a 128 x TBN tile is measured
// (8192^3 forced onto each tile, profiles/r01_tc_tiles.jsonl): 1 : 0.65 : 0.51 for TBN = 256 / 128 / 64 --
// the narrower the tile, the more its shared-memory operand reads ((128 + TBN) * 32 bytes per step)
// outweigh its tensor time (128 * TBN * 8 multiply-adds).
// The narrow tiles fill the GPU for the mid-size products (1024^3 is 32 tiles of 128 x 256 on 148 SMs)
// and trim the partial last wave of the large ones.
int pick_tc_tile(const TcPlan &t, int planes = 1)
{
    const int forced = g_tuning[ACCBLAS_TUNE_TC_TILE];
    if (forced >= 0 && forced <= 2) return forced;
    if (forced == 3) return t.M > TBM ? 3 : 0;                     // the pair tile needs two row blocks
    static const int tbn[3] = {256, 128, 64};
    static const double step_cost[3] = {134.0, 87.0, 68.0};
    const int sms = sm_count();
    const int64_t tm = (t.M + TBM - 1) / TBM;
    int best = 0;
    double best_cost = 0.0;
    for (int c = 0; c < 3; ++c) {
        const int64_t tiles = tm * ((t.N + tbn[c] - 1) / tbn[c]) * planes;
        // per tile: the K loop plus a prologue / epilogue worth ~40 K-steps of the widest tile
        const double cost = (double)((tiles + sms - 1) / sms) * (step_cost[c] * (double)(t.pitch / 8) + 40.0 * 134.0);
        if (c == 0 || cost < best_cost * 0.97) { best = c; best_cost = cost; }
    }
    // 3 = the 256 x 256 CTA-pair tile (cta_group::2): whenever the wide tile wins and pairing the row blocks does not
    // add a round (an odd count of 128-row blocks wastes half a pair in the last tile row)
    if (best == 0 && t.M > TBM) {
        const int64_t tn = (t.N + PBN - 1) / PBN;
        const int64_t rounds1 = (tm * tn * planes + sms - 1) / sms;
        const int64_t pairs = std::max(1, sms / 2), rounds2 = (((tm + 1) / 2) * tn * planes + pairs - 1) / pairs;
        if (rounds2 <= rounds1) best = 3;
    }
    return best;
}

// the part of SgemmParams that does not depend on the tile shape
void fill_tc_params(SgemmParams &p, const GemmArgs &g, const TcPlan &t, int planes, int plane_rows_a, int plane_rows_b,
                    int64_t plane_stride_c)
{
    p.M = t.M; p.N = t.N; p.Kp = t.pitch;      // padding columns are zero-filled by pack_split
    p.ldc = (g.dtype == ACCBLAS_C32 ? 2 : 1) * g.ldc;
    p.C = static_cast<float *>(g.C);
    p.cplx = g.dtype == ACCBLAS_C32;
    if (g.alpha_dev) p.alpha = Alpha<float2>{make_float2(0.f, 0.f), static_cast<const float2 *>(g.alpha)};
    else p.alpha = Alpha<float2>{make_float2(static_cast<const float *>(g.alpha)[0],
                                            p.cplx ? static_cast<const float *>(g.alpha)[1] : 0.f), nullptr};
    p.beta = BetaC<float>{nullptr, 0, 0.f};
    p.cbeta = make_float2(0.f, 0.f);
    if (planes == 1 && has_beta(g)) {                              // (the 3M planes take beta in combine3)
        p.beta.cin = static_cast<const float *>(g.Cin);
        p.beta.ld = (p.cplx ? 2 : 1) * (g.ldcin ? g.ldcin : g.ldc);
        p.beta.beta = static_cast<const float *>(g.beta)[0];
        if (p.cplx) p.cbeta = make_float2(static_cast<const float *>(g.beta)[0], static_cast<const float *>(g.beta)[1]);
    }
    p.plane_rows_a = plane_rows_a; p.plane_rows_b = plane_rows_b; p.plane_stride_c = plane_stride_c;
    p.refit = planes == 1 ? make_refit32(g) : no_refit32();      // the 3M planes are intermediates: combine3 refits
}

template <int TBN, int STAGES, int KCB>
int launch_tc(const GemmArgs &g, const TcPlan &t, float *ahi, float *alo, float *bhi, float *blo,
              int planes = 1, int plane_rows_a = 0, int plane_rows_b = 0, int64_t plane_stride_c = 0)
{
    using Cfg = TcCfg<TBN, STAGES>;
    CUtensorMap mAhi, mAlo, mBhi, mBlo;
    const uint64_t pitch_b = (uint64_t)t.pitch * 4;
    const CUtensorMapDataType F = CU_TENSOR_MAP_DATA_TYPE_FLOAT32;
    // stacked planes: the maps cover all of them (plane p starts at row p * plane_rows_*)
    const uint64_t rows_a = planes > 1 ? (uint64_t)planes * (uint64_t)plane_rows_a : (uint64_t)t.RA;
    const uint64_t rows_b = planes > 1 ? (uint64_t)planes * (uint64_t)plane_rows_b : (uint64_t)t.RB;
    if (!tma::make_2d(&mAhi, F, 4, ahi, rows_a, (uint64_t)t.pitch, TBM, TBK, pitch_b) ||
        !tma::make_2d(&mAlo, F, 4, alo, rows_a, (uint64_t)t.pitch, TBM, TBK, pitch_b) ||
        !tma::make_2d(&mBhi, F, 4, bhi, rows_b, (uint64_t)t.pitch, TBN, TBK, pitch_b) ||
        !tma::make_2d(&mBlo, F, 4, blo, rows_b, (uint64_t)t.pitch, TBN, TBK, pitch_b))
        return cuda_fail(cudaErrorInvalidValue, "cuTensorMapEncodeTiled");
    static thread_local AttrOnce attr_once;
    if (attr_once.need()) {
        ACC_CUDA(cudaFuncSetAttribute(sgemm_tf32x3_kernel<TBN, STAGES, KCB>, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg::SMEM));
    }
    SgemmParams p;
    fill_tc_params(p, g, t, planes, plane_rows_a, plane_rows_b, plane_stride_c);
    p.tiles_m = (int)((t.M + TBM - 1) / TBM);
    p.tiles_n = (int)((t.N + TBN - 1) / TBN);
    const int64_t tiles = (int64_t)p.tiles_m * p.tiles_n * planes;
    if (tiles > 0x7fffffff) return ACCBLAS_ERR_SHAPE;
    p.tiles_per_plane = planes > 1 ? p.tiles_m * p.tiles_n : 0;
    sgemm_tf32x3_kernel<TBN, STAGES, KCB><<<(unsigned)tiles, TTHREADS, Cfg::SMEM, g.stream>>>(mAhi, mAlo, mBhi, mBlo, p);
    return check_launch("sgemm_tf32x3_kernel");
}

// pair kernel (cta_group::2): pair tiles of 256 x 256, grid = 2 CTAs per tile launched as clusters of 2
template <int STAGES, int KCB>
int launch_tc_pair(const GemmArgs &g, const TcPlan &t, float *ahi, float *alo, float *bhi, float *blo,
                   int planes = 1, int plane_rows_a = 0, int plane_rows_b = 0, int64_t plane_stride_c = 0)
{
    using Cfg = PairCfg<STAGES>;
    CUtensorMap mAhi, mAlo, mBhi, mBlo;
    const uint64_t pitch_b = (uint64_t)t.pitch * 4;
    const CUtensorMapDataType F = CU_TENSOR_MAP_DATA_TYPE_FLOAT32;
    const uint64_t rows_a = planes > 1 ? (uint64_t)planes * (uint64_t)plane_rows_a : (uint64_t)t.RA;
    const uint64_t rows_b = planes > 1 ? (uint64_t)planes * (uint64_t)plane_rows_b : (uint64_t)t.RB;
    // every box is 128 rows: a CTA's rows of A, a CTA's HALF of the tile's B rows
    if (!tma::make_2d(&mAhi, F, 4, ahi, rows_a, (uint64_t)t.pitch, TBM, TBK, pitch_b) ||
        !tma::make_2d(&mAlo, F, 4, alo, rows_a, (uint64_t)t.pitch, TBM, TBK, pitch_b) ||
        !tma::make_2d(&mBhi, F, 4, bhi, rows_b, (uint64_t)t.pitch, PBN / 2, TBK, pitch_b) ||
        !tma::make_2d(&mBlo, F, 4, blo, rows_b, (uint64_t)t.pitch, PBN / 2, TBK, pitch_b))
        return cuda_fail(cudaErrorInvalidValue, "cuTensorMapEncodeTiled");
    static thread_local AttrOnce attr_once;
    if (attr_once.need()) {
        ACC_CUDA(cudaFuncSetAttribute(sgemm_tf32x3_pair_kernel<STAGES, KCB>, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg::SMEM));
    }
    SgemmParams p;
    fill_tc_params(p, g, t, planes, plane_rows_a, plane_rows_b, plane_stride_c);
    p.tiles_m = (int)((t.M + 2 * TBM - 1) / (2 * TBM));
    p.tiles_n = (int)((t.N + PBN - 1) / PBN);
    const int64_t tiles = (int64_t)p.tiles_m * p.tiles_n * planes;
    if (2 * tiles > 0x7fffffff) return ACCBLAS_ERR_SHAPE;
    p.tiles_per_plane = planes > 1 ? p.tiles_m * p.tiles_n : 0;
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = dim3((unsigned)(2 * tiles));
    cfg.blockDim = dim3(TTHREADS);
    cfg.dynamicSmemBytes = Cfg::SMEM;
    cfg.stream = g.stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = 2; attr[0].val.clusterDim.y = 1; attr[0].val.clusterDim.z = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    cudaError_t e = cudaLaunchKernelEx(&cfg, sgemm_tf32x3_pair_kernel<STAGES, KCB>, mAhi, mAlo, mBhi, mBlo, p);
    if (e != cudaSuccess) return cuda_fail(e, "sgemm_tf32x3_pair_kernel launch");
    return check_launch("sgemm_tf32x3_pair_kernel");
}

// ---- 3M for c32 (see pack_split3_kernel) ----
struct Tc3Plan {
    TcPlan t;              // one real m x n x k product
    int64_t ldt;           // row pitch of the T planes (floats)
    int64_t rows_a, rows_b; // padded rows of one A / B plane
    size_t plane_t;
};

Tc3Plan make_plan3(const GemmArgs &g)
{
    Tc3Plan q;
    q.t.RA = g.m; q.t.RB = g.n; q.t.Kp = g.k;
    q.t.pitch = (g.k + 31) / 32 * 32;
    q.t.M = g.m; q.t.N = g.n;
    // planes are stacked along the rows of one tensor map per operand: pad the row counts to the largest tile so a
    // TMA box never runs into the next plane's valid rows (rows past m / n only feed outputs that are not stored)
    q.rows_a = (g.m + 127) / 128 * 128;
    q.rows_b = (g.n + 255) / 256 * 256;
    q.t.plane_a = (size_t)q.rows_a * (size_t)q.t.pitch * 4;
    q.t.plane_b = (size_t)q.rows_b * (size_t)q.t.pitch * 4;
    q.ldt = (g.n + 3) / 4 * 4;
    q.plane_t = ((size_t)g.m * (size_t)q.ldt * 4 + 1023) / 1024 * 1024;
    return q;
}

int cgemm_3m_run(const GemmArgs &g)
{
    const Tc3Plan q = make_plan3(g);
    const size_t pa = q.t.plane_a, pb = q.t.plane_b, pt = q.plane_t;
    char *ws = nullptr;
    if (scratch_alloc(reinterpret_cast<void **>(&ws), 6 * (pa + pb) + 3 * pt, g.stream) != cudaSuccess) {
        (void)cudaGetLastError();                          // no room for 3M's planes: the 4M embedding needs a third of them
        return sgemm_tc(g);
    }
    char *wa = ws, *wb = ws + 6 * pa, *wt = wb + 6 * pb;
    Pack3Args a{}, b{};
    a.src = static_cast<const float2 *>(g.A); a.ld = g.lda;
    a.transposed = g.opA != ACCBLAS_OP_N;
    a.conj_sign = g.opA == ACCBLAS_OP_H ? -1.f : 1.f;
    a.R = g.m; a.Kp = g.k; a.pitch = q.t.pitch;
    b.src = static_cast<const float2 *>(g.B); b.ld = g.ldb;
    b.transposed = g.opB == ACCBLAS_OP_N;                  // K-major row j of op(B)^T is a stored column when op = N
    b.conj_sign = g.opB == ACCBLAS_OP_H ? -1.f : 1.f;
    b.R = g.n; b.Kp = g.k; b.pitch = q.t.pitch;
    for (int p = 0; p < 3; ++p) {
        a.hi[p] = reinterpret_cast<float *>(wa + p * pa); a.lo[p] = reinterpret_cast<float *>(wa + (3 + p) * pa);
        b.hi[p] = reinterpret_cast<float *>(wb + p * pb); b.lo[p] = reinterpret_cast<float *>(wb + (3 + p) * pb);
    }
    int s = ACCBLAS_OK;
    {
        dim3 grid((unsigned)((q.t.pitch + 31) / 32), (unsigned)((std::max(a.R, b.R) + 31) / 32), 2);
        if (grid.y > 65535u) s = ACCBLAS_ERR_SHAPE;
        else {
            pack_split3_kernel<<<grid, 256, 0, g.stream>>>(a, b);
            s = check_launch("pack_split3_kernel");
        }
    }
    static const float one = 1.f;
    const int tile = pick_tc_tile(q.t, 3);
    if (s == ACCBLAS_OK) {
        // ONE launch for the three products (planes stacked in the maps): 3 x tiles share the waves
        GemmArgs h = g;
        h.dtype = ACCBLAS_F32;
        h.alpha = &one; h.alpha_dev = false;
        h.beta = nullptr; h.Cin = nullptr;
        h.C = wt; h.ldc = q.ldt;
        const int ra = (int)q.rows_a, rb = (int)q.rows_b;
        const int64_t sc = (int64_t)(pt / 4);
        s = tile == 3 ? launch_tc_pair<3, 2>(h, q.t, a.hi[0], a.lo[0], b.hi[0], b.lo[0], 3, ra, rb, sc)
          : tile == 1 ? launch_tc<128, 3, 2>(h, q.t, a.hi[0], a.lo[0], b.hi[0], b.lo[0], 3, ra, rb, sc)
          : tile == 2 ? launch_tc<64, 4, 2>(h, q.t, a.hi[0], a.lo[0], b.hi[0], b.lo[0], 3, ra, rb, sc)
                      : launch_tc<256, 2, 2>(h, q.t, a.hi[0], a.lo[0], b.hi[0], b.lo[0], 3, ra, rb, sc);
    }
    if (s == ACCBLAS_OK) {
        Combine3Args c{};
        for (int p = 0; p < 3; ++p) c.t[p] = reinterpret_cast<const float *>(wt + p * pt);
        c.C = static_cast<float2 *>(g.C);
        c.m = g.m; c.n = g.n; c.ldt = q.ldt; c.ldc = g.ldc;
        c.alpha = make_alpha<float2>(g.alpha, g.alpha_dev);
        c.beta = make_beta<float2>(g);
        c.refit = make_refit32(g);
        const int64_t total = c.m * (c.ldt / 4);
        const int64_t want = (total + 255) / 256, cap = (int64_t)sm_count() * 32;
        combine3_kernel<<<(unsigned)std::max<int64_t>(1, std::min(want, cap)), 256, 0, g.stream>>>(c);
        s = check_launch("combine3_kernel");
    }
    cudaError_t e = cudaFreeAsync(ws, g.stream);
    if (s == ACCBLAS_OK && e != cudaSuccess) return cuda_fail(e, "cudaFreeAsync");
    return s;
}

}  // namespace

bool sgemm_tc_eligible(const GemmArgs &g)
{
    if (g.dtype != ACCBLAS_F32 && g.dtype != ACCBLAS_C32) return false;
    if (g.m < 1 || g.n < 1 || g.k < 1) return false;
    const TcPlan t = make_plan(g);
    if (t.RA > 2000000 || t.RB > 2000000 || t.Kp > 0x7fffffff) return false;
    // the pack pass and its workspace only pay off once the product is large enough
    if (t.M < 64 || t.N < 64 || t.Kp < 32) return false;
    if ((double)t.M * (double)t.N * (double)t.Kp < 16.0 * 1024 * 1024) return false;
    return true;
}

size_t sgemm_tc_workspace(const GemmArgs &g)
{
    const TcPlan t = make_plan(g);
    return 2 * (t.plane_a + t.plane_b);
}

int sgemm_tc(const GemmArgs &g)
{
    const TcPlan t = make_plan(g);
    const bool cx = g.dtype == ACCBLAS_C32;
    char *ws = nullptr;
    ACC_CUDA(scratch_alloc(reinterpret_cast<void **>(&ws), 2 * (t.plane_a + t.plane_b), g.stream));
    float *ahi = reinterpret_cast<float *>(ws), *alo = reinterpret_cast<float *>(ws + t.plane_a);
    float *bhi = reinterpret_cast<float *>(ws + 2 * t.plane_a), *blo = reinterpret_cast<float *>(ws + 2 * t.plane_a + t.plane_b);

    PackArgs pa{};
    pa.src = static_cast<const float *>(g.A); pa.rows = g.rowsA; pa.cols = g.colsA; pa.ld = g.lda;
    pa.transposed = g.opA != ACCBLAS_OP_N;                 // op(A)[i][l] lives at stored [l][i]
    pa.conj_sign = (cx && g.opA == ACCBLAS_OP_H) ? -1.f : 1.f;
    pa.R = t.RA; pa.Kp = t.Kp; pa.pitch = t.pitch; pa.hi = ahi; pa.lo = alo;
    PackArgs pb{};
    pb.src = static_cast<const float *>(g.B); pb.rows = g.rowsB; pb.cols = g.colsB; pb.ld = g.ldb;
    pb.transposed = g.opB == ACCBLAS_OP_N;                 // K-major row j of op(B)^T is a stored COLUMN when op = N
    pb.conj_sign = (cx && g.opB == ACCBLAS_OP_H) ? -1.f : 1.f;
    pb.R = t.RB; pb.Kp = t.Kp; pb.pitch = t.pitch; pb.hi = bhi; pb.lo = blo;

    int s = launch_pack(cx, pa, pb, g.stream);
    if (s == ACCBLAS_OK) {
        // K-blocks per accumulation chunk (ACCBLAS_TC_KCB).  Measured on the pair tile at 8192^3 (profiles/r02_error_table.jsonl,
        // r02_pair_kcb.jsonl): chunks of 2 / 4 / 8 give a componentwise error of 1.1 / 1.7 / 2.8 u (u = 2^-24; cublasSgemm on
        // the same inputs: 6.1 u) at the same speed within the power-cap noise (243 / 241 TFLOP/s) -- so the default stays 2.
        static const int kcb = [] { const char *e = getenv("ACCBLAS_TC_KCB"); return e ? atoi(e) : 2; }();
        const int tile = pick_tc_tile(t);
        s = tile == 3 ? (kcb == 4 ? launch_tc_pair<3, 4>(g, t, ahi, alo, bhi, blo)
                       : kcb == 8 ? launch_tc_pair<3, 8>(g, t, ahi, alo, bhi, blo)
                                  : launch_tc_pair<3, 2>(g, t, ahi, alo, bhi, blo))
          : tile == 1 ? launch_tc<128, 3, 2>(g, t, ahi, alo, bhi, blo)
          : tile == 2 ? launch_tc<64, 4, 2>(g, t, ahi, alo, bhi, blo)
          : kcb == 1 ? launch_tc<256, 2, 1>(g, t, ahi, alo, bhi, blo)
          : kcb == 4 ? launch_tc<256, 2, 4>(g, t, ahi, alo, bhi, blo)
          : kcb >= 1000 ? launch_tc<256, 2, 1 << 30>(g, t, ahi, alo, bhi, blo)
                     : launch_tc<256, 2, 2>(g, t, ahi, alo, bhi, blo);
    }
    cudaError_t e = cudaFreeAsync(ws, g.stream);
    if (s == ACCBLAS_OK && e != cudaSuccess) return cuda_fail(e, "cudaFreeAsync");
    return s;
}

size_t cgemm_3m_workspace(const GemmArgs &g)
{
    const Tc3Plan q = make_plan3(g);
    return 6 * (q.t.plane_a + q.t.plane_b) + 3 * q.plane_t;
}

// 3M pays once the tensor time saved (a quarter of the 4M product, at the ~250 TFLOP/s the kernel sustains)
// outweighs the extra streaming pass (combine + the wider T planes) and three launches instead of one.
bool cgemm_3m_wanted(const GemmArgs &g)
{
    const int knob = g_tuning[ACCBLAS_TUNE_CGEMM_3M];
    if (knob == 0 || g.dtype != ACCBLAS_C32 || !sgemm_tc_eligible(g)) return false;
    if (g.m < 128 || g.n < 64 || g.k < 32) return false;
    if (knob == 1) return true;
    const double save = 2.0 * (double)g.m * (double)g.n * (double)g.k / 250.0e12;
    const double cost = 32.0 * (double)g.m * (double)g.n / 4.0e12 + 10.0e-6;
    return save > cost;          // measured crossover ~1536^3 (profiles/r01_cgemm_3m.jsonl)
}

int cgemm_3m(const GemmArgs &g) { return cgemm_3m_run(g); }

}  // namespace accblas
