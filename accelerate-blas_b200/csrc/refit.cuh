// refit.cuh -- non-finite inputs in the split-precision (3xTF32) f32 / c32 tensor kernels.
//
// The split  a = a_hi + a_lo  cannot carry a non-finite value: a = +Inf gives a_hi = +Inf and a_lo = Inf - Inf = NaN,
// and even with a_lo forced to 0 the cross terms produce Inf * b_lo (of either sign, or Inf * 0) next to Inf * b_hi,
// so an output that cublasSgemm / cublasCgemm -- what the reference calls (LLVM/PTX/Level3.hs:91,101) -- returns
// as +-Inf came out as NaN (round-1 advisor finding).  A value whose TF32 rounding overflows (|a| within 2^-12 of
// FLT_MAX) has the same problem.
//
// Every such case makes the ACCUMULATED output non-finite, and a non-finite output is rare and cheap to detect in
// the epilogue.  So the kernels keep the fast split, and the warp that owns a non-finite accumulator recomputes
// that one inner product  op(A)[i,:] . op(B)[:,j]  in plain FP32 FMAs from the ORIGINAL operands (all 32 lanes share
// the k range, partial sums combined by shuffles).  IEEE arithmetic then yields exactly what a non-split FP32 gemm
// yields: +-Inf where the terms agree in sign, NaN for Inf - Inf or Inf * 0, NaN wherever an input is NaN.
// Cost when every output is finite: one exponent test per output and one vote per warp.
#pragma once
#include "common.cuh"

namespace accblas {

struct Refit32 {
    const float *A, *B;      // the caller's operands (float view); A == nullptr: no refit (intermediate planes)
    int64_t lda, ldb;        // row pitches in ELEMENTS (complex elements for c32)
    int64_t k;
    int ta, tb;              // 1: the operand is stored transposed w.r.t. op(.) (op != N)
    float sa, sb;            // -1: conjugate the operand (op == H on a complex type), else +1
    int64_t batch_a, batch_b;  // FLOATS between the operands of consecutive products of a strided batch (0: none / shared)
};

inline Refit32 make_refit32(const GemmArgs &g)
{
    Refit32 f;
    f.A = static_cast<const float *>(g.A);
    f.B = static_cast<const float *>(g.B);
    f.lda = g.lda; f.ldb = g.ldb; f.k = g.k;
    f.ta = g.opA != ACCBLAS_OP_N; f.tb = g.opB != ACCBLAS_OP_N;
    const bool cx = g.dtype == ACCBLAS_C32;
    f.sa = (cx && g.opA == ACCBLAS_OP_H) ? -1.f : 1.f;
    f.sb = (cx && g.opB == ACCBLAS_OP_H) ? -1.f : 1.f;
    f.batch_a = g.batch > 1 ? g.strideA * (cx ? 2 : 1) : 0;
    f.batch_b = g.batch > 1 ? g.strideB * (cx ? 2 : 1) : 0;
    return f;
}
inline Refit32 no_refit32() { Refit32 f{}; f.A = nullptr; return f; }

__device__ __forceinline__ bool nonfinite32(float v) { return (__float_as_uint(v) & 0x7f800000u) == 0x7f800000u; }

// whole warp, converged: sum_l op(A)[i][l] * op(B)[l][j] in FP32.  `a`, `b` point at the first term, sa / sb are the
// element strides along l.  (Scalars, not the struct: a by-reference struct argument of a noinline function costs the
// calling kernel a stack frame.)
static __device__ __noinline__ float refit_dot_real(const float *a, const float *b, int64_t sa, int64_t sb, int64_t k)
{
    float s = 0.f;
    for (int64_t l = threadIdx.x & 31; l < k; l += 32) s = fmaf(a[l * sa], b[l * sb], s);
    return warp_sum(s);
}
static __device__ __noinline__ float2 refit_dot_cplx(const float2 *a, const float2 *b, int64_t sa, int64_t sb, int64_t k,
                                              float conj_a, float conj_b)
{
    float2 s = make_float2(0.f, 0.f);
    for (int64_t l = threadIdx.x & 31; l < k; l += 32) {
        float2 x = a[l * sa], y = b[l * sb];
        x.y *= conj_a; y.y *= conj_b;
        s = fma_acc(x, y, s);
    }
    return warp_sum(s);
}

// Called by ALL 32 lanes of a converged warp, each with its own accumulator value and the (row, column) it belongs to
// (`mine` = the lane holds a real output there).  Returns the value to use.
__device__ __forceinline__ float refit_real(const Refit32 &f, float v, int64_t r, int64_t c, bool mine, int64_t batch = 0)
{
    unsigned bad = __ballot_sync(0xffffffffu, mine && nonfinite32(v));
    while (bad) {
        const int src = __ffs(bad) - 1;
        bad &= bad - 1;
        const int64_t i = __shfl_sync(0xffffffffu, r, src), j = __shfl_sync(0xffffffffu, c, src);
        const float s = refit_dot_real(f.A + batch * f.batch_a + (f.ta ? i : i * f.lda), f.B + batch * f.batch_b + (f.tb ? j * f.ldb : j),
                                       f.ta ? f.lda : 1, f.tb ? 1 : f.ldb, f.k);
        if ((int)(threadIdx.x & 31) == src) v = s;
    }
    return v;
}
__device__ __forceinline__ float2 refit_cplx(const Refit32 &f, float2 v, int64_t r, int64_t c, bool mine, int64_t batch = 0)
{
    unsigned bad = __ballot_sync(0xffffffffu, mine && (nonfinite32(v.x) || nonfinite32(v.y)));
    while (bad) {
        const int src = __ffs(bad) - 1;
        bad &= bad - 1;
        const int64_t i = __shfl_sync(0xffffffffu, r, src), j = __shfl_sync(0xffffffffu, c, src);
        const float2 s = refit_dot_cplx(reinterpret_cast<const float2 *>(f.A + batch * f.batch_a) + (f.ta ? i : i * f.lda),
                                        reinterpret_cast<const float2 *>(f.B + batch * f.batch_b) + (f.tb ? j * f.ldb : j),
                                        f.ta ? f.lda : 1, f.tb ? 1 : f.ldb, f.k, f.sa, f.sb);
        if ((int)(threadIdx.x & 31) == src) v = s;
    }
    return v;
}

}  // namespace accblas
