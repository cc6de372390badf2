/*
 * oracle/oracle.c -- TEST INFRASTRUCTURE ONLY.  Never linked into, imported by or
 * called from the product (libaccblas_b200.so / accblas_b200 python host).  Only
 * tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
 * legs may load the shared object built from this file.
 *
 * A plain-C CPU restatement of the reference's OWN definition of the hot path --
 * the pure-Accelerate fallbacks that the reference's test-suite evaluates with the
 * Accelerate interpreter and compares the BLAS hooks against
 * (test/Test/BLAS/Level3.hs:54, test/Test/BLAS/Level2.hs:50):
 *
 *   mXm   src/Data/Array/Accelerate/Numeric/LinearAlgebra/BLAS/Level3.hs:82-110
 *   mXv   src/Data/Array/Accelerate/Numeric/LinearAlgebra/BLAS/Level2.hs:79-93
 *   dotu  src/Data/Array/Accelerate/Numeric/LinearAlgebra/BLAS/Level1.hs:73-74
 *   dotc  src/Data/Array/Accelerate/Numeric/LinearAlgebra/BLAS/Level1.hs:85-92
 *   sdot  src/Data/Array/Accelerate/Numeric/LinearAlgebra/BLAS/Level1.hs:44-62
 *   asum / amax / amin   src/.../BLAS/Level1.hs:100-109, 115-126, 132-143
 *
 * shape rules follow the hooks:
 *   gemm  src/.../LLVM/PTX/Level3.hs:54-65   (m,k) = opA==N ? (rowsA,colsA) : (colsA,rowsA);
 *                                            n = opB==N ? colsB : rowsB
 *   gemv  src/.../LLVM/PTX/Level2.hs:78-87   |y| = opA==N ? rowsA : colsA
 *   dtype src/.../LinearAlgebra/Type.hs:32-36 complex = interleaved (re,im) pairs
 *   op    src/.../LinearAlgebra/Type.hs:78-90 H on a real type behaves as T
 *
 * Two evaluation modes:
 *   wide = 0  "literal": element-type arithmetic, left fold from 0 of (alpha*a)*b,
 *             exactly what the interpreter does with the fallback.
 *   wide = 1  "truth": the same sum accumulated one precision up (double for f32 /
 *             c32, long double for f64 / c64) and rounded once at the end; this is
 *             what normwise error bounds are measured against.
 *
 * Parity pin: the golden vectors are the haddock examples of the reference
 * (src/.../LinearAlgebra.hs:48-56, 65-71, 86-95, 106-115, 129-149), see
 * tests/golden/.  The reference itself is Haskell (no ghc/cabal/stack in this image,
 * no network) and cannot be built, so there is no oracle/_ref binary; the second pin
 * is the third-party library its native backend calls, OpenBLAS, driven with the
 * reference's exact CBLAS arguments from oracle/openblas_ref.py.
 * dotu/dotc have NO test in the reference (test/Test/BLAS.hs:29-32): beyond the two
 * haddock examples their parity is unpinned by the reference.
 */
#include <stdint.h>
#include <stddef.h>

enum { DT_F32 = 0, DT_F64 = 1, DT_C32 = 2, DT_C64 = 3 };
enum { OP_N = 0, OP_T = 1, OP_H = 2 };

/* element (i,j) of op(X) for a stored rows x cols row-major X (Type.hs:78-90) */
#define IDX(op, ld, i, j) ((op) == OP_N ? (size_t)(i) * (size_t)(ld) + (size_t)(j) \
                                        : (size_t)(j) * (size_t)(ld) + (size_t)(i))

/* ---- real kernels.  T = storage type, W = accumulation type.  With W == T the
 *      arithmetic is the literal element-type fold (compile with -ffp-contract=off
 *      so that no product is fused into the add). ---------------------------------- */
#define REAL_GEMM(NAME, T, W)                                                        \
static void NAME(int opA, int opB, int64_t m, int64_t n, int64_t k, int64_t lda,      \
                 int64_t ldb, T alpha, const T *A, const T *B, T *C)                 \
{                                                                                    \
    for (int64_t i = 0; i < m; ++i)                                                  \
        for (int64_t j = 0; j < n; ++j) {                                            \
            W acc = 0;                           /* fold (+) 0        Level3.hs:83 */\
            for (int64_t l = 0; l < k; ++l) {                                        \
                W a = (W)A[IDX(opA, lda, i, l)];                                     \
                W b = (W)B[IDX(opB, ldb, l, j)];                                     \
                W t = (W)alpha * a;              /* alpha * a * b     Level3.hs:84 */\
                W p = t * b;                                                         \
                acc = acc + p;                                                       \
            }                                                                        \
            C[(size_t)i * (size_t)n + (size_t)j] = (T)acc;                           \
        }                                                                            \
}

REAL_GEMM(gemm_f32_lit,  float,  float)
REAL_GEMM(gemm_f32_wide, float,  double)
REAL_GEMM(gemm_f64_lit,  double, double)
REAL_GEMM(gemm_f64_wide, double, long double)

/* ---- complex kernels: interleaved storage, conj folded into the read --------- */
#define CPLX_GEMM(NAME, T, W)                                                        \
static void NAME(int opA, int opB, int64_t m, int64_t n, int64_t k, int64_t lda,      \
                 int64_t ldb, const T *alpha, const T *A, const T *B, T *C)          \
{                                                                                    \
    const W sa = (opA == OP_H) ? (W)-1 : (W)1;   /* map conjugate  Level3.hs:98-99 */\
    const W sb = (opB == OP_H) ? (W)-1 : (W)1;   /*                Level3.hs:108-109*/\
    const W alr = (W)alpha[0], ali = (W)alpha[1];                                    \
    for (int64_t i = 0; i < m; ++i)                                                  \
        for (int64_t j = 0; j < n; ++j) {                                            \
            W accr = 0, acci = 0;                                                    \
            for (int64_t l = 0; l < k; ++l) {                                        \
                size_t ia = 2 * IDX(opA, lda, i, l), ib = 2 * IDX(opB, ldb, l, j);   \
                W ar = (W)A[ia], ai = sa * (W)A[ia + 1];                             \
                W br = (W)B[ib], bi = sb * (W)B[ib + 1];                             \
                W t1 = alr * ar, t2 = ali * ai, t3 = alr * ai, t4 = ali * ar;        \
                W tr = t1 - t2, ti = t3 + t4;                                        \
                W p1 = tr * br, p2 = ti * bi, p3 = tr * bi, p4 = ti * br;            \
                W pr = p1 - p2, pi = p3 + p4;                                        \
                accr = accr + pr; acci = acci + pi;                                  \
            }                                                                        \
            size_t ic = 2 * ((size_t)i * (size_t)n + (size_t)j);                     \
            C[ic] = (T)accr; C[ic + 1] = (T)acci;                                    \
        }                                                                            \
}

CPLX_GEMM(gemm_c32_lit,  float,  float)
CPLX_GEMM(gemm_c32_wide, float,  double)
CPLX_GEMM(gemm_c64_lit,  double, double)
CPLX_GEMM(gemm_c64_wide, double, long double)

static int bad_dtype(int dt) { return dt < 0 || dt > 3; }
static int bad_op(int op) { return op < 0 || op > 2; }

/*
 * C (m x n, row-major, tight) = alpha * op(A) * op(B).   Returns 0, or 1 on a bad
 * argument (the reference validates nothing: PTX/Level3.hs:54-65 never compares the
 * inner extents; the restatement refuses a mismatch instead of reading out of bounds).
 */
int oracle_gemm(int dtype, int wide, int opA, int opB,
                int64_t rowsA, int64_t colsA, int64_t rowsB, int64_t colsB,
                const void *alpha, const void *A, const void *B, void *C)
{
    if (bad_dtype(dtype) || bad_op(opA) || bad_op(opB)) return 1;
    if (rowsA < 0 || colsA < 0 || rowsB < 0 || colsB < 0) return 1;
    int64_t m  = opA == OP_N ? rowsA : colsA;      /* PTX/Level3.hs:57-59 */
    int64_t k  = opA == OP_N ? colsA : rowsA;
    int64_t n  = opB == OP_N ? colsB : rowsB;      /* PTX/Level3.hs:60-62 */
    int64_t kb = opB == OP_N ? rowsB : colsB;
    if (k != kb) return 1;
    int64_t lda = colsA, ldb = colsB;              /* PTX/Level3.hs:64-65 */
    int ra = opA, rb = opB;
    switch (dtype) {
    case DT_F32:
        if (ra == OP_H) ra = OP_T;                 /* Level3.hs:100 */
        if (rb == OP_H) rb = OP_T;                 /* Level3.hs:110 */
        (wide ? gemm_f32_wide : gemm_f32_lit)(ra, rb, m, n, k, lda, ldb,
            *(const float *)alpha, (const float *)A, (const float *)B, (float *)C);
        break;
    case DT_F64:
        if (ra == OP_H) ra = OP_T;
        if (rb == OP_H) rb = OP_T;
        (wide ? gemm_f64_wide : gemm_f64_lit)(ra, rb, m, n, k, lda, ldb,
            *(const double *)alpha, (const double *)A, (const double *)B, (double *)C);
        break;
    case DT_C32:
        (wide ? gemm_c32_wide : gemm_c32_lit)(opA, opB, m, n, k, lda, ldb,
            (const float *)alpha, (const float *)A, (const float *)B, (float *)C);
        break;
    default:
        (wide ? gemm_c64_wide : gemm_c64_lit)(opA, opB, m, n, k, lda, ldb,
            (const double *)alpha, (const double *)A, (const double *)B, (double *)C);
        break;
    }
    return 0;
}

/*
 * y = alpha * op(A) * x   (mXv, Level2.hs:79-93).  It is the n = 1 column case of
 * mXm with opB = N, which is also literally how the reference routes the complex-H
 * case on the GPU (PTX/Level2.hs:57-69, x viewed as a (len x 1) matrix).
 */
int oracle_gemv(int dtype, int wide, int opA, int64_t rowsA, int64_t colsA,
                const void *alpha, const void *A, const void *x, int64_t lenx, void *y)
{
    if (bad_dtype(dtype) || bad_op(opA) || lenx < 0) return 1;
    int64_t need = opA == OP_N ? colsA : rowsA;
    if (lenx != need) return 1;
    return oracle_gemm(dtype, wide, opA, OP_N, rowsA, colsA, lenx, 1, alpha, A, x, y);
}

/*
 * out = sum_i x_i * y_i            (dotu, Level1.hs:73-74)
 * out = sum_i conj(x_i) * y_i      (dotc, Level1.hs:85-92; conj on the FIRST vector)
 * The caller passes n = min(|x|,|y|): zipWith intersects the extents (Level1.hs:74).
 */
int oracle_dot(int dtype, int wide, int conj_x, int64_t n,
               const void *x, const void *y, void *out)
{
    if (bad_dtype(dtype) || n < 0) return 1;
    switch (dtype) {
    case DT_F32: {
        const float *a = x, *b = y;
        if (wide) { double s = 0; for (int64_t i = 0; i < n; ++i) s += (double)a[i] * (double)b[i];
                    *(float *)out = (float)s; }
        else      { float s = 0;  for (int64_t i = 0; i < n; ++i) s = s + a[i] * b[i];
                    *(float *)out = s; }
        break; }
    case DT_F64: {
        const double *a = x, *b = y;
        if (wide) { long double s = 0; for (int64_t i = 0; i < n; ++i) s += (long double)a[i] * (long double)b[i];
                    *(double *)out = (double)s; }
        else      { double s = 0; for (int64_t i = 0; i < n; ++i) s = s + a[i] * b[i];
                    *(double *)out = s; }
        break; }
    case DT_C32: {
        const float *a = x, *b = y; float *o = out;
        const double sg = conj_x ? -1.0 : 1.0;
        if (wide) {
            double sr = 0, si = 0;
            for (int64_t i = 0; i < n; ++i) {
                double ar = a[2*i], ai = sg * a[2*i+1], br = b[2*i], bi = b[2*i+1];
                sr += ar * br - ai * bi; si += ar * bi + ai * br;
            }
            o[0] = (float)sr; o[1] = (float)si;
        } else {
            float sr = 0, si = 0, sf = (float)sg;
            for (int64_t i = 0; i < n; ++i) {
                float ar = a[2*i], ai = sf * a[2*i+1], br = b[2*i], bi = b[2*i+1];
                float p1 = ar * br, p2 = ai * bi, p3 = ar * bi, p4 = ai * br;
                float pr = p1 - p2, pi = p3 + p4;
                sr = sr + pr; si = si + pi;
            }
            o[0] = sr; o[1] = si;
        }
        break; }
    default: {
        const double *a = x, *b = y; double *o = out;
        const long double sg = conj_x ? -1.0L : 1.0L;
        if (wide) {
            long double sr = 0, si = 0;
            for (int64_t i = 0; i < n; ++i) {
                long double ar = a[2*i], ai = sg * a[2*i+1], br = b[2*i], bi = b[2*i+1];
                sr += ar * br - ai * bi; si += ar * bi + ai * br;
            }
            o[0] = (double)sr; o[1] = (double)si;
        } else {
            double sr = 0, si = 0, sd = (double)sg;
            for (int64_t i = 0; i < n; ++i) {
                double ar = a[2*i], ai = sd * a[2*i+1], br = b[2*i], bi = b[2*i+1];
                double p1 = ar * br, p2 = ai * bi, p3 = ar * bi, p4 = ai * br;
                double pr = p1 - p2, pi = p3 + p4;
                sr = sr + pr; si = si + pi;
            }
            o[0] = sr; o[1] = si;
        }
        break; }
    }
    return 0;
}

/*
 * sdot z xs ys (Level1.hs:44-62): every element is converted to double (complex double), the
 * products are summed by a left fold STARTING FROM z (`fold (+) z' (zipWith (*) xs' ys')`), and the
 * result is converted back to the element type.  literal (wide = 0) is that double-precision fold;
 * wide = 1 carries the sum in long double (the truth error bounds are measured against).
 */
int oracle_sdot(int dtype, int wide, int64_t n, const void *z, const void *x, const void *y, void *out)
{
    if (bad_dtype(dtype) || n < 0) return 1;
    const int cplx = dtype == DT_C32 || dtype == DT_C64;
    const int f32 = dtype == DT_F32 || dtype == DT_C32;
#define LD(p, i) (f32 ? (double)((const float *)(p))[i] : ((const double *)(p))[i])
    if (!cplx) {
        if (wide) { long double s = LD(z, 0); for (int64_t i = 0; i < n; ++i) s += (long double)LD(x, i) * (long double)LD(y, i);
                    if (f32) *(float *)out = (float)(double)s; else *(double *)out = (double)s; }
        else      { double s = LD(z, 0); for (int64_t i = 0; i < n; ++i) s = s + LD(x, i) * LD(y, i);
                    if (f32) *(float *)out = (float)s; else *(double *)out = s; }
    } else {
        if (wide) {
            long double sr = LD(z, 0), si = LD(z, 1);
            for (int64_t i = 0; i < n; ++i) {
                long double ar = LD(x, 2*i), ai = LD(x, 2*i+1), br = LD(y, 2*i), bi = LD(y, 2*i+1);
                sr += ar * br - ai * bi; si += ar * bi + ai * br;
            }
            if (f32) { ((float *)out)[0] = (float)(double)sr; ((float *)out)[1] = (float)(double)si; }
            else     { ((double *)out)[0] = (double)sr; ((double *)out)[1] = (double)si; }
        } else {
            double sr = LD(z, 0), si = LD(z, 1);
            for (int64_t i = 0; i < n; ++i) {
                double ar = LD(x, 2*i), ai = LD(x, 2*i+1), br = LD(y, 2*i), bi = LD(y, 2*i+1);
                double p1 = ar * br, p2 = ai * bi, p3 = ar * bi, p4 = ai * br;
                sr = sr + (p1 - p2); si = si + (p3 + p4);
            }
            if (f32) { ((float *)out)[0] = (float)sr; ((float *)out)[1] = (float)si; }
            else     { ((double *)out)[0] = sr; ((double *)out)[1] = si; }
        }
    }
    return 0;
}

/*
 * asum (Level1.hs:100-109): `sum . map abs`, complex `mag c = abs (real c) + abs (imag c)`; the
 * result has the element's base type.  literal = element-precision left fold; wide = long double.
 */
int oracle_asum(int dtype, int wide, int64_t n, const void *x, void *out)
{
    if (bad_dtype(dtype) || n < 0) return 1;
    const int cplx = dtype == DT_C32 || dtype == DT_C64;
    const int f32 = dtype == DT_F32 || dtype == DT_C32;
    if (f32) {
        const float *a = x;
        if (wide) { long double s = 0; for (int64_t i = 0; i < n; ++i) s += cplx ? (long double)__builtin_fabsf(a[2*i]) + __builtin_fabsf(a[2*i+1]) : (long double)__builtin_fabsf(a[i]);
                    *(float *)out = (float)(double)s; }
        else      { float s = 0; for (int64_t i = 0; i < n; ++i) { float m = cplx ? __builtin_fabsf(a[2*i]) + __builtin_fabsf(a[2*i+1]) : __builtin_fabsf(a[i]); s = s + m; }
                    *(float *)out = s; }
    } else {
        const double *a = x;
        if (wide) { long double s = 0; for (int64_t i = 0; i < n; ++i) s += cplx ? (long double)__builtin_fabs(a[2*i]) + __builtin_fabs(a[2*i+1]) : (long double)__builtin_fabs(a[i]);
                    *(double *)out = (double)s; }
        else      { double s = 0; for (int64_t i = 0; i < n; ++i) { double m = cplx ? __builtin_fabs(a[2*i]) + __builtin_fabs(a[2*i+1]) : __builtin_fabs(a[i]); s = s + m; }
                    *(double *)out = s; }
    }
    return 0;
}

/*
 * amax / amin (Level1.hs:115-143): `fold1 cmp . indexed . map mag` with
 *   cmp ix iy = snd ix > snd iy ? ix : iy        (amax;  `<` for amin)
 * restated as the sequential left fold: the accumulator is ix, the next element iy, so the next
 * element replaces the accumulator unless the accumulator is STRICTLY better -- on ties the later
 * index wins.  Magnitudes are computed in the element's own precision.  Returns the 0-based index
 * in *out (int64, Haskell Int); n == 0 is an error (fold1 of an empty array).
 */
int oracle_iamax(int dtype, int want_max, int64_t n, const void *x, int64_t *out)
{
    if (bad_dtype(dtype) || n <= 0) return 1;
    const int cplx = dtype == DT_C32 || dtype == DT_C64;
    int64_t best = 0;
    if (dtype == DT_F32 || dtype == DT_C32) {
        const float *a = x;
        volatile float bv = cplx ? __builtin_fabsf(a[0]) + __builtin_fabsf(a[1]) : __builtin_fabsf(a[0]);
        for (int64_t i = 1; i < n; ++i) {
            volatile float v = cplx ? __builtin_fabsf(a[2*i]) + __builtin_fabsf(a[2*i+1]) : __builtin_fabsf(a[i]);
            const int keep = want_max ? (bv > v) : (bv < v);
            if (!keep) { bv = v; best = i; }
        }
    } else {
        const double *a = x;
        volatile double bv = cplx ? __builtin_fabs(a[0]) + __builtin_fabs(a[1]) : __builtin_fabs(a[0]);
        for (int64_t i = 1; i < n; ++i) {
            volatile double v = cplx ? __builtin_fabs(a[2*i]) + __builtin_fabs(a[2*i+1]) : __builtin_fabs(a[i]);
            const int keep = want_max ? (bv > v) : (bv < v);
            if (!keep) { bv = v; best = i; }
        }
    }
    *out = best;
    return 0;
}

int oracle_abi_version(void) { return 2; }
