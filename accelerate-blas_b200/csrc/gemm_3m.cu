// gemm_3m.cu -- large complex-double products as THREE real DMMA products ("3M"), the other half of
// the 4M/3M choice for BLAS.zgemm (LLVM/PTX/Level3.hs:106).
//
//   op(A) = Ar + i Ai,  op(B) = Br + i Bi   (the conjugation of `H` is a sign on Ai / Bi)
//   T1 = Ar Br,   T2 = Ai Bi,   T3 = (Ar + Ai)(Br + Bi)
//   C  = alpha * ((T1 - T2) + i (T3 - T1 - T2))
//
// 6 mnk real multiply-adds instead of the 8 mnk of the in-register 4M kernel (gemm_dmma.cu), paid
// for with three O(n^2) streaming passes:
//   1. split_planes_kernel  : interleaved (re,im) operand, as stored -> planes re, s*im, re + s*im
//                             (s = -1 under H).  The planes keep the STORED shape: the transpose is
//                             left to the real DMMA kernel, which consumes either major directly.
//   2. dgemm_dmma x 3       : the real f64 kernel, one launch per plane pair, alpha = 1.
//   3. combine_planes_kernel: T1, T2, T3 -> interleaved C, times the complex alpha.  C is only written.
// The FP64 tensor pipe is the bound (36.9 TFLOP/s): at 4096^3 the three products take 3 x 3.85 ms
// against 15.3 ms for 4M; the passes move 2.0 GB (~0.4 ms).
//
// Numerics: normwise the 3M error obeys the same k*eps*|A||B| bound as 4M (tests use the stated
// complex tolerance); component-wise Im C is formed by cancellation of T3 against T1 + T2, so an
// imaginary part that is tiny next to |A||B| loses relative accuracy.  The automatic choice therefore
// only takes 3M for large products where the saving is real (cost model below), and
// accblas_set_tuning(ACCBLAS_TUNE_ZGEMM_3M, 0) pins 4M.
#include "common.cuh"
#include <algorithm>

namespace accblas {
namespace {

// 16-byte streaming store (planes and C are written once and read by a later kernel from L2/HBM)
__device__ __forceinline__ void st2(double *p, double a, double b) {
    *reinterpret_cast<double2 *>(p) = make_double2(a, b);
}

struct SplitArgs {
    const double2 *src;      // rows x cols complex, row pitch ld (complex elements)
    double *p0, *p1, *p2;    // planes rows x cols, row pitch ldp (doubles, even)
    int64_t rows, cols, ld, ldp;
    int64_t quads_per_row;   // ceil(cols / 4)
    double sign;             // +1, or -1 for a conjugated operand
};

// One thread = four consecutive complex elements of a row: 64 contiguous bytes in, 32 contiguous
// bytes out per plane.  Grid-stride over (row, quad).
__global__ void __launch_bounds__(256) split_planes_kernel(const SplitArgs a)
{
    const int64_t total = a.rows * a.quads_per_row;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
        const int64_t r = i / a.quads_per_row, c = (i - r * a.quads_per_row) * 4;
        const double2 *s = a.src + r * a.ld + c;
        double2 z[4];
#pragma unroll
        for (int e = 0; e < 4; ++e) {
            if (c + e < a.cols) {
                const int4 raw = ldg_stream(s + e);
                z[e].x = __hiloint2double(raw.y, raw.x);
                z[e].y = __hiloint2double(raw.w, raw.z) * a.sign;
            } else {
                z[e] = make_double2(0.0, 0.0);
            }
        }
        const int64_t o = r * a.ldp + c;
#pragma unroll
        for (int h = 0; h < 2; ++h) {
            if (c + 2 * h < a.cols) {                        // ldp is even: a started pair always fits the row
                st2(a.p0 + o + 2 * h, z[2 * h].x, z[2 * h + 1].x);
                st2(a.p1 + o + 2 * h, z[2 * h].y, z[2 * h + 1].y);
                st2(a.p2 + o + 2 * h, z[2 * h].x + z[2 * h].y, z[2 * h + 1].x + z[2 * h + 1].y);
            }
        }
    }
}

struct CombineArgs {
    const double *t1, *t2, *t3;   // m x n, row pitch ldt (even)
    double2 *C;                   // m x n complex, row pitch ldc
    int64_t m, n, ldt, ldc, pairs_per_row;
    Alpha<double2> alpha;
    BetaC<double2> beta;
};

__global__ void __launch_bounds__(256) combine_planes_kernel(const CombineArgs a)
{
    const double2 al = a.alpha.get();
    const int64_t total = a.m * a.pairs_per_row;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
        const int64_t r = i / a.pairs_per_row, c = (i - r * a.pairs_per_row) * 2;
        const int64_t o = r * a.ldt + c;
        const int4 r1 = ldg_stream(a.t1 + o), r2 = ldg_stream(a.t2 + o), r3 = ldg_stream(a.t3 + o);
        const double t1[2] = {__hiloint2double(r1.y, r1.x), __hiloint2double(r1.w, r1.z)};
        const double t2[2] = {__hiloint2double(r2.y, r2.x), __hiloint2double(r2.w, r2.z)};
        const double t3[2] = {__hiloint2double(r3.y, r3.x), __hiloint2double(r3.w, r3.z)};
        double2 *out = a.C + r * a.ldc + c;
#pragma unroll
        for (int e = 0; e < 2; ++e) {
            if (c + e < a.n) {
                const double xr = t1[e] - t2[e], xi = (t3[e] - t1[e]) - t2[e];
                double2 v = make_double2(al.x * xr - al.y * xi, al.x * xi + al.y * xr);
                if (a.beta.on()) v = a.beta.add(v, r, c + e);
                out[e] = v;
            }
        }
    }
}

int64_t even_up(int64_t v) { return (v + 1) & ~(int64_t)1; }
size_t plane_bytes(int64_t rows, int64_t cols) { return ((size_t)rows * (size_t)even_up(cols) * 8 + 255) / 256 * 256; }

int launch_split(cudaStream_t st, const void *src, int64_t rows, int64_t cols, int64_t ld, bool conj, char *planes)
{
    SplitArgs a;
    a.src = static_cast<const double2 *>(src);
    const size_t pb = plane_bytes(rows, cols);
    a.p0 = reinterpret_cast<double *>(planes);
    a.p1 = reinterpret_cast<double *>(planes + pb);
    a.p2 = reinterpret_cast<double *>(planes + 2 * pb);
    a.rows = rows; a.cols = cols; a.ld = ld; a.ldp = even_up(cols);
    a.quads_per_row = (cols + 3) / 4;
    a.sign = conj ? -1.0 : 1.0;
    const int64_t total = rows * a.quads_per_row;
    const int64_t want = (total + 255) / 256, cap = (int64_t)sm_count() * 32;
    split_planes_kernel<<<(unsigned)std::max<int64_t>(1, std::min(want, cap)), 256, 0, st>>>(a);
    return check_launch("split_planes_kernel");
}

}  // namespace

size_t zgemm_3m_workspace(const GemmArgs &g)
{
    return 3 * (plane_bytes(g.rowsA, g.colsA) + plane_bytes(g.rowsB, g.colsB) + plane_bytes(g.m, g.n));
}

// Is 3M possible at all for this product (alignment, real-kernel eligibility of the plane products)?
bool zgemm_3m_eligible(const GemmArgs &g)
{
    if (g.dtype != ACCBLAS_C64) return false;
    if (g.m < 1 || g.n < 1 || g.k < 1) return false;
    if ((reinterpret_cast<uintptr_t>(g.A) | reinterpret_cast<uintptr_t>(g.B) | reinterpret_cast<uintptr_t>(g.C)) & 15u) return false;
    if (g.rowsA > 0x7fffffff || g.colsA > 0x7fffffff || g.rowsB > 0x7fffffff || g.colsB > 0x7fffffff) return false;
    return g.m * g.n >= 64 * 64 && g.k >= 8;            // what dgemm_dmma_eligible asks of the plane products
}

// Automatic choice: the saving is a quarter of the 4M tensor time (2 mnk of its 8 mnk flop at the DMMA
// rate); the price is the three streaming passes (40 B per stored operand element, 64 B per element
// of C counting the extra plane stores of the real kernel) plus five more launches.
bool zgemm_3m_wanted(const GemmArgs &g)
{
    const int knob = g_tuning[ACCBLAS_TUNE_ZGEMM_3M];
    if (knob == 0 || !zgemm_3m_eligible(g)) return false;
    if (knob == 1) return true;
    const double save = 2.0 * (double)g.m * (double)g.n * (double)g.k / 36.0e12;          // (8 - 6) mnk flop
    const double cost = (40.0 * ((double)g.rowsA * (double)g.colsA + (double)g.rowsB * (double)g.colsB) +
                         64.0 * (double)g.m * (double)g.n) / 4.0e12 + 12.0e-6;
    return save > 1.5 * cost;
}

int zgemm_3m(const GemmArgs &g)
{
    const size_t pa = plane_bytes(g.rowsA, g.colsA), pb = plane_bytes(g.rowsB, g.colsB), pt = plane_bytes(g.m, g.n);
    char *ws = nullptr;
    if (scratch_alloc(reinterpret_cast<void **>(&ws), 3 * (pa + pb + pt), g.stream) != cudaSuccess) {
        (void)cudaGetLastError();                          // no room for the planes: the in-register 4M kernel needs none
        return dgemm_dmma(g);
    }
    char *wa = ws, *wb = ws + 3 * pa, *wt = wb + 3 * pb;
    int st = launch_split(g.stream, g.A, g.rowsA, g.colsA, g.lda, g.opA == ACCBLAS_OP_H, wa);
    if (st == ACCBLAS_OK) st = launch_split(g.stream, g.B, g.rowsB, g.colsB, g.ldb, g.opB == ACCBLAS_OP_H, wb);
    static const double one = 1.0;
    for (int p = 0; p < 3 && st == ACCBLAS_OK; ++p) {
        GemmArgs h = g;
        h.dtype = ACCBLAS_F64;
        h.opA = g.opA == ACCBLAS_OP_N ? ACCBLAS_OP_N : ACCBLAS_OP_T;
        h.opB = g.opB == ACCBLAS_OP_N ? ACCBLAS_OP_N : ACCBLAS_OP_T;
        h.lda = even_up(g.colsA); h.ldb = even_up(g.colsB); h.ldc = even_up(g.n);
        h.alpha = &one; h.alpha_dev = false;
        h.beta = nullptr; h.Cin = nullptr;
        h.A = wa + p * pa; h.B = wb + p * pb; h.C = wt + p * pt;
        st = dgemm_dmma_eligible(h) ? dgemm_dmma(h) : ACCBLAS_ERR_SHAPE;
    }
    if (st == ACCBLAS_OK) {
        CombineArgs c;
        c.t1 = reinterpret_cast<const double *>(wt);
        c.t2 = reinterpret_cast<const double *>(wt + pt);
        c.t3 = reinterpret_cast<const double *>(wt + 2 * pt);
        c.C = static_cast<double2 *>(g.C);
        c.m = g.m; c.n = g.n; c.ldt = even_up(g.n); c.ldc = g.ldc; c.pairs_per_row = (g.n + 1) / 2;
        c.alpha = make_alpha<double2>(g.alpha, g.alpha_dev);
        c.beta = make_beta<double2>(g);
        const int64_t total = c.m * c.pairs_per_row;
        const int64_t want = (total + 255) / 256, cap = (int64_t)sm_count() * 32;
        combine_planes_kernel<<<(unsigned)std::max<int64_t>(1, std::min(want, cap)), 256, 0, g.stream>>>(c);
        st = check_launch("combine_planes_kernel");
    }
    const cudaError_t e = cudaFreeAsync(ws, g.stream);
    if (st == ACCBLAS_OK && e != cudaSuccess) return cuda_fail(e, "cudaFreeAsync");
    return st;
}

}  // namespace accblas
