// gemv.cu -- y = alpha * op(A) * x on row-major A (HBM-bound: the matrix is read once).
//
// Replaces BLAS.{s,d,c,z}gemv in as_gemv (LLVM/PTX/Level2.hs:105-124) AND the gemm detour
// the reference takes for complex `H` (LLVM/PTX/Level2.hs:52-53,57-69): working on the
// row-major data directly, conj-transpose is just another column-sum kernel.
//
//  op = N  (y_i = sum_j A[i][j] x_j; `#>`, LinearAlgebra.hs:99-101)
//      A persistent grid (CTAs resident = occupancy x SM count) walks units of ROWS rows.  All
//      threads of the CTA stride along the row in 16-byte vectors: ROWS streaming 128-bit
//      loads of A plus one cached 128-bit load of x in flight per thread, x re-used from
//      registers across the ROWS rows; warp-shuffle + shared-memory row reduction.
//  op = T / H  (y_j = sum_i op(A[i][j]) x_i; `<#`, LinearAlgebra.hs:119-121)
//      Column-owner threads: each thread owns one 16-byte column vector and marches down
//      the rows with 8 independent streaming loads in flight; x is staged in shared memory.
//      Rows are split over gridDim.y so that the whole grid is ONE resident wave; the
//      per-split partial rows are summed in split order by a second tiny kernel
//      (deterministic, no atomics).
#include "common.cuh"
#include <cstdlib>

namespace accblas {
namespace {

template <typename T> union Vec16 {
    int4 raw;
    T e[16 / sizeof(T)];
};

// ------------------------------------------------------------------ op = N ------------
template <typename T> struct GnRows { static constexpr int value = sizeof(T) == 16 ? 4 : 8; };
constexpr int GN_MAX_ROWS = 8;
constexpr int GN_MAX_WARPS = 16;

// NR rows starting at `row0` (all valid): every thread strides along the rows in 16-byte vectors.  The
// row addresses are row0 + r * n with r a compile-time constant -- no per-row offset registers (an
// earlier version kept a clamped offset per row and spilled under the 64-register cap of 2 x 512
// threads per SM, which cost 12 % of the bandwidth).
template <typename T, bool ALIGNED, int NR>
__device__ __forceinline__ void gemv_n_rows(int64_t n, int64_t nv, T alpha, const T *__restrict__ row0,
                                            const T *__restrict__ x, T *__restrict__ yout,
                                            T (*red)[GN_MAX_WARPS])
{
    constexpr int VEC = ALIGNED ? 16 / sizeof(T) : 1;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarps = blockDim.x >> 5;
    T acc[NR];
#pragma unroll
    for (int r = 0; r < NR; ++r) acc[r] = Scalar<T>::zero();
    if (ALIGNED) {
        for (int64_t v = threadIdx.x; v < nv; v += blockDim.x) {
            Vec16<T> xv, av[NR];
#pragma unroll
            for (int r = 0; r < NR; ++r)
                av[r].raw = ldg_stream(reinterpret_cast<const int4 *>(row0 + r * n) + v);
            xv.raw = ldg_cached(reinterpret_cast<const int4 *>(x) + v);
#pragma unroll
            for (int r = 0; r < NR; ++r)
#pragma unroll
                for (int e = 0; e < VEC; ++e) acc[r] = fma_acc(av[r].e[e], xv.e[e], acc[r]);
        }
    } else {
        for (int64_t j = threadIdx.x; j < n; j += blockDim.x) {
            const T xj = x[j];
#pragma unroll
            for (int r = 0; r < NR; ++r) acc[r] = fma_acc(row0[r * n + j], xj, acc[r]);
        }
    }
#pragma unroll
    for (int r = 0; r < NR; ++r) {
        acc[r] = warp_sum(acc[r]);
        if (lane == 0) red[r][warp] = acc[r];
    }
    __syncthreads();
    if (threadIdx.x < NR) {
        T s = Scalar<T>::zero();
        for (int w = 0; w < nwarps; ++w) s = add_s(s, red[threadIdx.x][w]);
        yout[threadIdx.x] = mul_s(alpha, s);
    }
    __syncthreads();
}

template <typename T, bool ALIGNED>
__global__ void __launch_bounds__(512, 2)
gemv_n_kernel(int64_t m, int64_t n, Alpha<T> alpha_, const T *__restrict__ A, const T *__restrict__ x,
              T *__restrict__ y)
{
    const T alpha = alpha_.get();
    constexpr int VEC = ALIGNED ? 16 / sizeof(T) : 1;
    constexpr int GN_ROWS = GnRows<T>::value;
    __shared__ T red[GN_ROWS][GN_MAX_WARPS];
    const int64_t nfull = m / GN_ROWS;
    const int64_t nv = n / VEC;
    for (int64_t u = blockIdx.x; u < nfull; u += gridDim.x)
        gemv_n_rows<T, ALIGNED, GN_ROWS>(n, nv, alpha, A + u * GN_ROWS * n, x, y + u * GN_ROWS, red);
    // the (at most GN_ROWS - 1) rows past the last full unit, one at a time, spread over the CTAs
    for (int64_t r = nfull * GN_ROWS + blockIdx.x; r < m; r += gridDim.x)
        gemv_n_rows<T, ALIGNED, 1>(n, nv, alpha, A + r * n, x, y + r, red);
}

// ------------------------------------------------------------------ op = T / H --------
constexpr int GT_THREADS = 256;
constexpr int GT_UNROLL = 8;
constexpr int GT_XCHUNK = 1024;     // rows of x staged per shared-memory refill

template <typename T, bool CONJ, bool ALIGNED>
__global__ void __launch_bounds__(GT_THREADS, 3)
gemv_t_kernel(int64_t m, int64_t n, int64_t rows_per_split, Alpha<T> alpha_, const T *__restrict__ A,
              const T *__restrict__ x, T *__restrict__ out /* y, or partial[split][n] */, bool direct)
{
    const T alpha = alpha_.get();
    constexpr int VEC = ALIGNED ? 16 / sizeof(T) : 1;
    __shared__ T xs[GT_XCHUNK];
    const int64_t nv = (n + VEC - 1) / VEC;    // ALIGNED guarantees n % VEC == 0
    const int64_t v = (int64_t)blockIdx.x * GT_THREADS + threadIdx.x;
    const bool active = v < nv;
    const int64_t r_begin = (int64_t)blockIdx.y * rows_per_split;
    int64_t r_end = r_begin + rows_per_split;
    if (r_end > m) r_end = m;

    T acc[VEC];
#pragma unroll
    for (int e = 0; e < VEC; ++e) acc[e] = Scalar<T>::zero();
    const char *colp = reinterpret_cast<const char *>(A) + (active ? v : 0) * (int64_t)(VEC * sizeof(T));
    const int64_t pitch = n * (int64_t)sizeof(T);

    for (int64_t rc = r_begin; rc < r_end; rc += GT_XCHUNK) {
        const int cnt = (int)((r_end - rc) < GT_XCHUNK ? (r_end - rc) : GT_XCHUNK);
        __syncthreads();
        for (int i = threadIdx.x; i < cnt; i += GT_THREADS) xs[i] = x[rc + i];
        __syncthreads();
        if (!active) continue;
        const char *p = colp + rc * pitch;
        int i = 0;
        for (; i + GT_UNROLL <= cnt; i += GT_UNROLL) {
            Vec16<T> av[GT_UNROLL];
#pragma unroll
            for (int u = 0; u < GT_UNROLL; ++u) {
                if (ALIGNED) av[u].raw = ldg_stream(p + (int64_t)(i + u) * pitch);
                else av[u].e[0] = *reinterpret_cast<const T *>(p + (int64_t)(i + u) * pitch);
            }
#pragma unroll
            for (int u = 0; u < GT_UNROLL; ++u) {
                const T xi = xs[i + u];
#pragma unroll
                for (int e = 0; e < VEC; ++e) acc[e] = fma_acc(CONJ ? conj_s(av[u].e[e]) : av[u].e[e], xi, acc[e]);
            }
        }
        for (; i < cnt; ++i) {
            Vec16<T> av;
            if (ALIGNED) av.raw = ldg_stream(p + (int64_t)i * pitch);
            else av.e[0] = *reinterpret_cast<const T *>(p + (int64_t)i * pitch);
            const T xi = xs[i];
#pragma unroll
            for (int e = 0; e < VEC; ++e) acc[e] = fma_acc(CONJ ? conj_s(av.e[e]) : av.e[e], xi, acc[e]);
        }
    }
    if (!active) return;
    T *dst = direct ? out : out + (int64_t)blockIdx.y * n;
#pragma unroll
    for (int e = 0; e < VEC; ++e) {
        const int64_t j = v * VEC + e;
        if (j < n) dst[j] = direct ? mul_s(alpha, acc[e]) : acc[e];
    }
}

// Small and medium matrices (the reference's own benchmark shapes, 200 x 400 .. 2000 x 3000,
// bench/Bench/Accelerate.hs:53-58): ONE launch, no scratch.  A CTA owns a slab of SLAB_CV 16-byte
// column vectors (128 contiguous bytes of every row) and splits the rows over SLAB_RG row groups;
// a warp load touches 4 rows x 128 bytes = four whole lines.  The row groups are summed through
// shared memory in a fixed order (deterministic).  Parallelism is columns / SLAB_CV CTAs, which is
// enough to fill the GPU exactly where the two-pass kernel above is launch-bound.
//
// Row split over a thread-block CLUSTER (round 2).  columns / SLAB_CV CTAs is a poor count at the top of the range:
// 2000 x 3000 doubles are 188 slabs on 148 SMs -- 40 SMs stream two slabs while 108 stream one, 63 % of the
// bandwidth at best (measured 0.85x cuBLAS, zgemv 0.75x).  A cluster of R = 2 / 4 / 8 CTAs now shares a slab, CTA r
// taking rows [r m / R, (r + 1) m / R); the R partial slabs are handed to the cluster's rank 0 through distributed
// shared memory and summed there in rank order -- still one launch, no scratch, no atomics, bit-reproducible --
// and the grid is R times finer (1504 CTAs for the example: 92 % balanced).  R is chosen per shape by slab_cluster().
constexpr int SLAB_CV = 8, SLAB_RG = 32, SLAB_UNROLL = 8, SLAB_MAXR = 8;

__device__ __forceinline__ uint32_t slab_cluster_rank() { uint32_t r; asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r)); return r; }
__device__ __forceinline__ uint32_t slab_cluster_size() { uint32_t r; asm volatile("mov.u32 %0, %%cluster_nctarank;" : "=r"(r)); return r; }

template <typename T, bool CONJ, bool ALIGNED>
__global__ void __launch_bounds__(SLAB_CV * SLAB_RG)
gemv_t_slab_kernel(int64_t m_all, int64_t n, Alpha<T> alpha_, const T *__restrict__ A, const T *__restrict__ x_all,
                   T *__restrict__ y)
{
    const T alpha = alpha_.get();
    constexpr int VEC = ALIGNED ? 16 / sizeof(T) : 1;
    __shared__ T red[SLAB_RG][SLAB_CV * VEC + 1];
    __shared__ T part[SLAB_MAXR][SLAB_CV * VEC];                  // rank 0: the other ranks' partial slabs
    const uint32_t R = slab_cluster_size(), rank = slab_cluster_rank();
    const int64_t slab = blockIdx.x / R;
    // "I am running": distributed shared memory of a CTA may only be touched once that CTA has started.  Arrive here,
    // wait just before the remote stores below -- the streaming loop in between hides the barrier's latency.
    if (R > 1) asm volatile("barrier.cluster.arrive.relaxed.aligned;" ::: "memory");
    // this CTA's rows [row0, row0 + m)
    const int64_t row0 = m_all * rank / R, m = m_all * (rank + 1) / R - row0;
    const T *__restrict__ x = x_all + row0;
    const int cv = threadIdx.x % SLAB_CV, rg = threadIdx.x / SLAB_CV;
    const int64_t nv = (n + VEC - 1) / VEC;
    const int64_t v = slab * SLAB_CV + cv;
    const bool active = v < nv;
    T acc[VEC];
#pragma unroll
    for (int e = 0; e < VEC; ++e) acc[e] = Scalar<T>::zero();
    if (active) {
        const int64_t pitch = n * (int64_t)sizeof(T);
        const char *p = reinterpret_cast<const char *>(A) + row0 * pitch + v * (int64_t)(VEC * sizeof(T));
        int64_t r = rg;
        for (; r + (int64_t)(SLAB_UNROLL - 1) * SLAB_RG < m; r += (int64_t)SLAB_UNROLL * SLAB_RG) {
            Vec16<T> av[SLAB_UNROLL];
            T xi[SLAB_UNROLL];
#pragma unroll
            for (int u = 0; u < SLAB_UNROLL; ++u) {
                const int64_t rr = r + (int64_t)u * SLAB_RG;
                if (ALIGNED) av[u].raw = ldg_stream(p + rr * pitch);
                else av[u].e[0] = *reinterpret_cast<const T *>(p + rr * pitch);
                xi[u] = __ldg(x + rr);
            }
#pragma unroll
            for (int u = 0; u < SLAB_UNROLL; ++u)
#pragma unroll
                for (int e = 0; e < VEC; ++e) acc[e] = fma_acc(CONJ ? conj_s(av[u].e[e]) : av[u].e[e], xi[u], acc[e]);
        }
        for (; r < m; r += SLAB_RG) {
            Vec16<T> av;
            if (ALIGNED) av.raw = ldg_stream(p + r * pitch);
            else av.e[0] = *reinterpret_cast<const T *>(p + r * pitch);
            const T xi = __ldg(x + r);
#pragma unroll
            for (int e = 0; e < VEC; ++e) acc[e] = fma_acc(CONJ ? conj_s(av.e[e]) : av.e[e], xi, acc[e]);
        }
    }
#pragma unroll
    for (int e = 0; e < VEC; ++e) red[rg][cv * VEC + e] = acc[e];
    __syncthreads();
    T s = Scalar<T>::zero();
    if (threadIdx.x < SLAB_CV * VEC) {
#pragma unroll 8
        for (int g = 0; g < SLAB_RG; ++g) s = add_s(s, red[g][threadIdx.x]);
    }
    if (R > 1) {
        // hand the partial slab to rank 0 (distributed shared memory), which adds them in rank order
        asm volatile("barrier.cluster.wait.aligned;" ::: "memory");      // every CTA of the cluster has started (see entry)
        if (rank != 0 && threadIdx.x < SLAB_CV * VEC) {
            const uint32_t local = (uint32_t)__cvta_generic_to_shared(&part[rank][threadIdx.x]);
            uint32_t remote;
            asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(remote) : "r"(local), "r"(0));
            if constexpr (sizeof(T) == 4) asm volatile("st.shared::cluster.b32 [%0], %1;" ::"r"(remote), "r"(*reinterpret_cast<const uint32_t *>(&s)) : "memory");
            else if constexpr (sizeof(T) == 8) asm volatile("st.shared::cluster.b64 [%0], %1;" ::"r"(remote), "l"(*reinterpret_cast<const uint64_t *>(&s)) : "memory");
            else asm volatile("st.shared::cluster.v2.b64 [%0], {%1, %2};" ::"r"(remote), "l"(reinterpret_cast<const uint64_t *>(&s)[0]),
                              "l"(reinterpret_cast<const uint64_t *>(&s)[1]) : "memory");
        }
        asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
        asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
        if (rank != 0) return;
        if (threadIdx.x < SLAB_CV * VEC)
            for (uint32_t q = 1; q < R; ++q) s = add_s(s, part[q][threadIdx.x]);
    }
    if (threadIdx.x < SLAB_CV * VEC) {
        const int64_t j = slab * SLAB_CV * VEC + threadIdx.x;
        if (j < n) y[j] = mul_s(alpha, s);
    }
}

// cluster size of the slab kernel: the R in {1, 2, 4, 8} whose grid of slabs * R CTAs loads the SMs most evenly
// (CTAs are small: several are resident per SM, so the balance is total / (ceil(total / SMs) * SMs)), the smaller
// R on near-ties; every CTA keeps at least two full unrolled passes of rows.  ACCBLAS_GEMV_T_CLUSTER=R forces it.
inline int slab_cluster(int64_t slabs, int64_t rows)
{
    static const int forced = [] { const char *e = getenv("ACCBLAS_GEMV_T_CLUSTER"); return e ? atoi(e) : 0; }();
    if (forced == 1 || forced == 2 || forced == 4 || forced == 8) return forced;      // (8: measurement only)
    const int64_t sms = sm_count();
    int best = 1;
    double best_eff = 0.0;
    // clusters of 8 measured SLOWER than no split (2000 x 3000 doubles: 14.3 us against 10.3 us for R = 2 or 4 and 12.3 us
    // for R = 1, profiles/r02_gemv_t_cluster.txt): placement of an 8-CTA cluster needs 8 free slots in one GPC
    // Fewer than two CTAs per SM: the kernel is bound by the bytes in flight per SM, not by balance -- split the rows once
    // (twice when even that leaves SMs empty).  Kernel times at 1000 x 2000 (ncu, us; R = 1 / 2 / 4): Float 5.6 / 5.0 / 5.1,
    // Double 5.9 / 5.2 / 5.4, Complex Float 7.2 / 5.9 / 6.1, Complex Double 6.7 / 6.2 / 9.6 (cuBLAS 5.7 / 6.3 / 6.4 / 7.9);
    // the balance rule below had picked 1, 1, 1 and 4 for them.
    const int64_t min_rows = 2 * SLAB_RG * SLAB_UNROLL / 4;
    if (slabs < 2 * sms) {
        if (rows / 2 < min_rows) return 1;
        return (2 * slabs < sms && rows / 4 >= min_rows) ? 4 : 2;
    }
    for (int R = 1; R <= 4; R *= 2) {
        if (R > 1 && rows / R < min_rows) break;
        const int64_t total = slabs * R;
        const double eff = (double)total / (double)(((total + sms - 1) / sms) * sms);
        if (eff > best_eff * 1.04) { best = R; best_eff = eff; }
    }
    return best;
}

template <typename T>
__global__ void __launch_bounds__(256)
gemv_t_reduce_kernel(int64_t n, int splits, Alpha<T> alpha_, const T *__restrict__ partial, T *__restrict__ y)
{
    const T alpha = alpha_.get();
    const int64_t j = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= n) return;
    T s = Scalar<T>::zero();
    for (int sp = 0; sp < splits; ++sp) s = add_s(s, partial[(int64_t)sp * n + j]);
    y[j] = mul_s(alpha, s);
}

template <typename T> __global__ void fill_zero_kernel(int64_t n, T *__restrict__ y)
{
    const int64_t j = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (j < n) y[j] = Scalar<T>::zero();
}

template <typename K> int occupancy(K kernel, int threads)
{
    int nb = 0;
    if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&nb, kernel, threads, 0) != cudaSuccess || nb < 1) nb = 1;
    return nb;
}

template <typename T>
int gemv_typed(cudaStream_t st, int opA, int64_t rowsA, int64_t colsA, const void *alpha_p, bool alpha_dev,
               const void *A_, const void *x_, void *y_)
{
    const Alpha<T> alpha = make_alpha<T>(alpha_p, alpha_dev);
    const T *A = static_cast<const T *>(A_), *x = static_cast<const T *>(x_);
    T *y = static_cast<T *>(y_);
    const int64_t leny = opA == ACCBLAS_OP_N ? rowsA : colsA;
    const int64_t lenx = opA == ACCBLAS_OP_N ? colsA : rowsA;
    if (leny == 0) return ACCBLAS_OK;
    if (lenx == 0) {   // empty sum: y = 0
        fill_zero_kernel<T><<<(unsigned)((leny + 255) / 256), 256, 0, st>>>(leny, y);
        return check_launch("fill_zero_kernel");
    }
    constexpr int VEC = 16 / sizeof(T);
    const bool aligned = (reinterpret_cast<uintptr_t>(A) & 15u) == 0 && (colsA % VEC) == 0;

    if (opA == ACCBLAS_OP_N) {
        const bool al = aligned && (reinterpret_cast<uintptr_t>(x) & 15u) == 0;
        const int64_t nv = al ? colsA / VEC : colsA;
        int threads = nv >= 512 ? 512 : nv >= 256 ? 256 : nv >= 128 ? 128 : 64;
        const int64_t units = (rowsA + GnRows<T>::value - 1) / GnRows<T>::value;
        static thread_local int occ_a = 0, occ_u = 0;
        if (!occ_a) { occ_a = occupancy(gemv_n_kernel<T, true>, 512); occ_u = occupancy(gemv_n_kernel<T, false>, 512); }
        int64_t grid = (int64_t)sm_count() * (al ? occ_a : occ_u) * (512 / threads);
        if (grid > units) grid = units;
        else grid = (units + (units + grid - 1) / grid - 1) / ((units + grid - 1) / grid);   // same rounds, even load
        if (al) gemv_n_kernel<T, true><<<(unsigned)grid, threads, 0, st>>>(rowsA, colsA, alpha, A, x, y);
        else    gemv_n_kernel<T, false><<<(unsigned)grid, threads, 0, st>>>(rowsA, colsA, alpha, A, x, y);
        return check_launch("gemv_n_kernel");
    }

    const bool conj = Scalar<T>::cplx && opA == ACCBLAS_OP_H;
    const int64_t nv = aligned ? colsA / VEC : colsA;
    static const int64_t slab_max = [] { const char *e = getenv("ACCBLAS_GEMV_T_SLAB_MAX"); return e ? atoll(e) : (int64_t)128 << 20; }();
    if ((int64_t)sizeof(T) * rowsA * colsA <= slab_max) {       // one launch, no scratch
        const int64_t slabs = (nv + SLAB_CV - 1) / SLAB_CV;
        // 64 .. 128 MiB (Complex Double 2000 x 3000 = 96 MB): one plain launch.  Kernel times there: slab R = 1 18.0 us, R = 2
        // 22.5, R = 4 21.1, the two-pass kernels 17.3 + 8.1 (cuBLAS 14.2).
        const int R = (int64_t)sizeof(T) * rowsA * colsA > ((int64_t)64 << 20) ? 1 : slab_cluster(slabs, rowsA);
        cudaLaunchConfig_t cfg{};
        cfg.gridDim = dim3((unsigned)(slabs * R));
        cfg.blockDim = dim3(SLAB_CV * SLAB_RG);
        cfg.stream = st;
        cudaLaunchAttribute attr[1];
        attr[0].id = cudaLaunchAttributeClusterDimension;
        attr[0].val.clusterDim.x = (unsigned)R; attr[0].val.clusterDim.y = 1; attr[0].val.clusterDim.z = 1;
        cfg.attrs = attr;
        cfg.numAttrs = 1;
        cudaError_t le = cudaSuccess;
        // R == 1 keeps the plain launch: a cluster launch costs ~2 us more even for clusters of one CTA (measured)
#define GO(CJ, AL) do { if (R == 1) gemv_t_slab_kernel<T, CJ, AL><<<cfg.gridDim, cfg.blockDim, 0, st>>>(rowsA, colsA, alpha, A, x, y); \
                        else le = cudaLaunchKernelEx(&cfg, gemv_t_slab_kernel<T, CJ, AL>, rowsA, colsA, alpha, A, x, y); } while (0)
        if (aligned) { if (conj) GO(true, true); else GO(false, true); }
        else         { if (conj) GO(true, false); else GO(false, false); }
#undef GO
        if (le != cudaSuccess) return cuda_fail(le, "gemv_t_slab_kernel launch");
        return check_launch("gemv_t_slab_kernel");
    }
    const int64_t gx = (nv + GT_THREADS - 1) / GT_THREADS;
    static thread_local int occ_t = 0;
    if (!occ_t) occ_t = occupancy(gemv_t_kernel<T, false, true>, GT_THREADS);
    const int64_t slots = (int64_t)sm_count() * occ_t;      // one resident wave
    int64_t splits = slots / gx;
    const int64_t max_splits = (rowsA + 63) / 64;            // at least 64 rows per split
    if (splits > max_splits) splits = max_splits;
    if (splits < 1) splits = 1;
    if (splits > 65535) splits = 65535;
    int64_t rps = (rowsA + splits - 1) / splits;
    rps = (rps + GT_UNROLL - 1) / GT_UNROLL * GT_UNROLL;
    splits = (rowsA + rps - 1) / rps;
    const bool direct = splits == 1;
    T *partial = nullptr;
    if (!direct) ACC_CUDA(scratch_alloc(reinterpret_cast<void **>(&partial), sizeof(T) * (size_t)splits * (size_t)colsA, st));
    dim3 grid((unsigned)gx, (unsigned)splits);
    T *out = direct ? y : partial;
#define GO(CJ, AL) gemv_t_kernel<T, CJ, AL><<<grid, GT_THREADS, 0, st>>>(rowsA, colsA, rps, alpha, A, x, out, direct)
    if (aligned) { if (conj) GO(true, true); else GO(false, true); }
    else         { if (conj) GO(true, false); else GO(false, false); }
#undef GO
    int s = check_launch("gemv_t_kernel");
    if (!direct) {
        if (s == ACCBLAS_OK) {
            gemv_t_reduce_kernel<T><<<(unsigned)((colsA + 255) / 256), 256, 0, st>>>(colsA, (int)splits, alpha, partial, y);
            s = check_launch("gemv_t_reduce_kernel");
        }
        cudaError_t e = cudaFreeAsync(partial, st);
        if (s == ACCBLAS_OK && e != cudaSuccess) return cuda_fail(e, "cudaFreeAsync");
    }
    return s;
}

}  // namespace

int gemv_dispatch(cudaStream_t s, int dtype, int opA, int64_t rowsA, int64_t colsA, const void *alpha, bool alpha_dev,
                  const void *A, const void *x, void *y)
{
    switch (dtype) {
    case ACCBLAS_F32: return gemv_typed<float>(s, opA, rowsA, colsA, alpha, alpha_dev, A, x, y);
    case ACCBLAS_F64: return gemv_typed<double>(s, opA, rowsA, colsA, alpha, alpha_dev, A, x, y);
    case ACCBLAS_C32: return gemv_typed<float2>(s, opA, rowsA, colsA, alpha, alpha_dev, A, x, y);
    default:          return gemv_typed<double2>(s, opA, rowsA, colsA, alpha, alpha_dev, A, x, y);
    }
}

}  // namespace accblas
