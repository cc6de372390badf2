// tests/cpp/hpp_check.cpp -- exercises include/accblas_b200.hpp (the C++ mirror of the reference's operator
// interface) on cuda:0 the way the reference's own suite does: the haddock known answers
// (LinearAlgebra.hs:48-56, 86-95, 106-115, 129-149) and random products over the four element types and the
// N / T / H flags (test/Test/BLAS/Level3.hs:31-72, Level2.hs:31-68) against a naive long-double restatement of
// mXm / mXv / dotu (BLAS/Level3.hs:82-110) written here -- test infrastructure, not a product path.
// Prints "hpp_check ok" and exits 0; any failure prints the case and exits 1.
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <random>
#include "../../include/accblas_b200.hpp"

using namespace accblas;
using cld = std::complex<long double>;
static int failures = 0;
#define EXPECT(cond, ...) do { if (!(cond)) { ++failures; printf("FAIL %s:%d: ", __FILE__, __LINE__); printf(__VA_ARGS__); printf("\n"); } } while (0)

template <class T> struct is_cplx { static constexpr bool value = false; };
template <class R> struct is_cplx<std::complex<R>> { static constexpr bool value = true; };
template <class T> cld widen(T v) { if constexpr (is_cplx<T>::value) return cld(v.real(), v.imag()); else return cld(v, 0); }
template <class T> T narrow(cld v) { if constexpr (is_cplx<T>::value) return T((typename T::value_type)v.real(), (typename T::value_type)v.imag()); else return (T)v.real(); }
template <class T> T rnd(std::mt19937_64 &g) {
    std::uniform_real_distribution<double> u(-1.0, 1.0);                   // test/Test/Util.hs:34-35
    if constexpr (is_cplx<T>::value) return T((typename T::value_type)u(g), (typename T::value_type)u(g));
    else return (T)u(g);
}
template <class T> long double eps_of() { return std::ldexp(1.0L, std::is_same<typename numeric_traits<T>::real, float>::value ? -24 : -53); }

// op(X)[i][l] of a stored rows x cols row-major matrix
template <class T> cld at(const std::vector<T> &x, int64_t cols, Transpose op, int64_t i, int64_t l) {
    if (op == Transpose::N) return widen(x[i * cols + l]);
    cld v = widen(x[l * cols + i]);
    return op == Transpose::H ? std::conj(v) : v;
}

template <class T> void gemm_case(std::mt19937_64 &g, Transpose oa, Transpose ob, int64_t m, int64_t n, int64_t k) {
    const int64_t ra = oa == Transpose::N ? m : k, ca = oa == Transpose::N ? k : m;
    const int64_t rb = ob == Transpose::N ? k : n, cb = ob == Transpose::N ? n : k;
    std::vector<T> A((size_t)(ra * ca)), B((size_t)(rb * cb));
    for (auto &v : A) v = rnd<T>(g);
    for (auto &v : B) v = rnd<T>(g);
    const T alpha = rnd<T>(g);
    Matrix<T> dA(ra, ca, A), dB(rb, cb, B);
    const std::vector<T> C = gemm(alpha, oa, dA, ob, dB).to_host();
    long double err2 = 0, na = 0, nb = 0;
    for (auto &v : A) na += std::norm(widen(v));
    for (auto &v : B) nb += std::norm(widen(v));
    for (int64_t i = 0; i < m; ++i)
        for (int64_t j = 0; j < n; ++j) {
            cld s = 0;
            for (int64_t l = 0; l < k; ++l) s += (widen(alpha) * at(A, ca, oa, i, l)) * at(B, cb, ob, l, j);
            err2 += std::norm(widen(C[i * n + j]) - s);
        }
    const long double bound = 2.0L * (k + 2) * eps_of<T>() * std::abs(widen(alpha)) * std::sqrt(na) * std::sqrt(nb);
    EXPECT(std::sqrt(err2) <= bound, "gemm dtype %d op %d%d %ldx%ldx%ld err %Lg bound %Lg", numeric_traits<T>::dtype, (int)oa, (int)ob,
           (long)m, (long)n, (long)k, std::sqrt(err2), bound);
}

template <class T> void gemv_dot_case(std::mt19937_64 &g, Transpose oa, int64_t rows, int64_t cols) {
    std::vector<T> A((size_t)(rows * cols)), x((size_t)(oa == Transpose::N ? cols : rows));
    for (auto &v : A) v = rnd<T>(g);
    for (auto &v : x) v = rnd<T>(g);
    const T alpha = rnd<T>(g);
    Matrix<T> dA(rows, cols, A);
    Vector<T> dx(x);
    const std::vector<T> y = gemv(alpha, oa, dA, dx).to_host();
    const int64_t leny = oa == Transpose::N ? rows : cols, lenx = (int64_t)x.size();
    long double err2 = 0, na = 0, nx = 0;
    for (auto &v : A) na += std::norm(widen(v));
    for (auto &v : x) nx += std::norm(widen(v));
    EXPECT((int64_t)y.size() == leny, "gemv length");
    for (int64_t i = 0; i < leny; ++i) {
        cld s = 0;
        for (int64_t l = 0; l < lenx; ++l) s += (widen(alpha) * at(A, cols, oa, i, l)) * widen(x[l]);
        err2 += std::norm(widen(y[i]) - s);
    }
    const long double bound = 2.0L * (lenx + 2) * eps_of<T>() * std::abs(widen(alpha)) * std::sqrt(na) * std::sqrt(nx);
    EXPECT(std::sqrt(err2) <= bound, "gemv dtype %d op %d %ldx%ld err %Lg bound %Lg", numeric_traits<T>::dtype, (int)oa, (long)rows, (long)cols,
           std::sqrt(err2), bound);
    // dotu / dotc on x with itself shifted: conj applies to the FIRST argument (BLAS/Level1.hs:85-92)
    std::vector<T> z(x.rbegin(), x.rend());
    Vector<T> dz(z);
    cld su = 0, sc = 0;
    for (int64_t l = 0; l < lenx; ++l) { su += widen(x[l]) * widen(z[l]); sc += std::conj(widen(x[l])) * widen(z[l]); }
    const long double db = (lenx + 2) * eps_of<T>() * nx;
    EXPECT(std::abs(widen(dotu(dx, dz).value()) - su) <= db, "dotu dtype %d n %ld", numeric_traits<T>::dtype, (long)lenx);
    EXPECT(std::abs(widen(dotc(dx, dz).value()) - sc) <= db, "dotc dtype %d n %ld", numeric_traits<T>::dtype, (long)lenx);
    // the compensated dot is at least as close (twice-double evaluation, one rounding to the element type)
    EXPECT(std::abs(widen(dot_compensated(dx, dz, false).value()) - su) <= 4 * eps_of<T>() * std::abs(su) + 1e-30L,
           "dot_compensated dtype %d n %ld", numeric_traits<T>::dtype, (long)lenx);
}

static void haddock() {
    using cd = std::complex<double>;
    // LinearAlgebra.hs:48-56  (<.>)
    EXPECT(dot(Vector<double>({1, 2, 3, 4}), Vector<double>({-2, 0, 1, 1})).value() == 5.0, "haddock dotu real");
    EXPECT(dot(Vector<cd>({cd(1, 1), cd(1, 0)}), Vector<cd>({cd(1, 0), cd(1, -1)})).value() == cd(2, 0), "haddock dotu complex");
    // Numeric/Sum.hs:72-80  sum kbn [1.0, 1.0e100, 1.0, -1.0e100] == 2.0  (as a dot with ones)
    EXPECT(dot_compensated(Vector<double>({1.0, 1.0e100, 1.0, -1.0e100}), Vector<double>({1, 1, 1, 1})).value() == 2.0, "haddock sum kbn");
    // LinearAlgebra.hs:65-71  (><)
    EXPECT(outer(Vector<double>({1, 2, 3}), Vector<double>({5, 2, 3})).to_host() == std::vector<double>({5, 2, 3, 10, 4, 6, 15, 6, 9}), "haddock outer");
    // LinearAlgebra.hs:86-95  (#>)   and :106-115  (<#)
    Matrix<double> m(2, 3, {1, 2, 3, 4, 5, 6});
    EXPECT(mXv(m, Vector<double>({10, 20, 30})).to_host() == std::vector<double>({140, 320}), "haddock mXv");
    EXPECT(vXm(Vector<double>({5, 10}), m).to_host() == std::vector<double>({45, 60, 75}), "haddock vXm");
    // LinearAlgebra.hs:129-149  (<>)
    Matrix<double> a(3, 5, {1, 2, 3, 4, 5, 6, 7, 8, 9, 10, 11, 12, 13, 14, 15}), b(5, 2, {1, 3, 0, 2, -1, 5, 7, 7, 6, 0});
    EXPECT(mXm(a, b).to_host() == std::vector<double>({56, 50, 121, 135, 186, 220}), "haddock mXm");
}

static void level1_and_errors() {
    // sdot / asum / amax / amin (BLAS/Level1.hs:44-139): small exact cases
    Vector<float> x({1.f, -7.f, 3.f, 7.f}), y({2.f, 1.f, 1.f, 1.f});
    EXPECT(sdot(0.5f, x, y).value() == 5.5f, "sdot");
    EXPECT(asum(x).value() == 18.f, "asum");
    EXPECT(amax(x).value() == 3, "amax returns the LAST index among ties (BLAS/Level1.hs:124)");
    EXPECT(amin(x).value() == 0, "amin");
    using cf = std::complex<float>;
    EXPECT(asum(Vector<cf>({cf(1, -2), cf(-3, 4)})).value() == 10.f, "asum complex = sum |re| + |im|");
    // the C ABI validates what the reference does not (inner extents): surfaces as accblas::Error
    bool threw = false;
    try { mXm(Matrix<double>(2, 3), Matrix<double>(4, 2)); } catch (const Error &e) { threw = e.status() == ACCBLAS_ERR_SHAPE; }
    EXPECT(threw, "inner-extent mismatch must raise ACCBLAS_ERR_SHAPE");
    // zero extents: k = 0 gives zeros (alpha * empty sum), empty outputs are fine
    const std::vector<double> z = mXm(Matrix<double>(3, 0), Matrix<double>(0, 2)).to_host();
    EXPECT(z == std::vector<double>(6, 0.0), "k = 0 must zero-fill");
    EXPECT(mXm(Matrix<double>(0, 4), Matrix<double>(4, 3)).rows() == 0, "m = 0");
}

template <class T> void sweep(std::mt19937_64 &g) {
    const Transpose ops[3] = {Transpose::N, Transpose::T, Transpose::H};
    std::uniform_int_distribution<int> dim(1, 64), dim2(1, 128);       // the reference's generators
    for (Transpose oa : ops)
        for (Transpose ob : ops) {
            gemm_case<T>(g, oa, ob, dim(g), dim(g), dim(g));
            gemm_case<T>(g, oa, ob, 200, 136, 72);                          // tensor-path tiles with ragged edges
        }
    gemm_case<T>(g, Transpose::N, Transpose::N, 300, 260, 1100);
    gemm_case<T>(g, Transpose::N, Transpose::N, 37, 41, 1);                 // k = 1: the outer-product kernel
    for (Transpose oa : ops) {
        gemv_dot_case<T>(g, oa, dim2(g), dim2(g));
        gemv_dot_case<T>(g, oa, 777, 1300);
    }
}

int main() {
    try {
        std::mt19937_64 g(20260104);
        haddock();
        level1_and_errors();
        sweep<float>(g);
        sweep<double>(g);
        sweep<std::complex<float>>(g);
        sweep<std::complex<double>>(g);
        // a second stream: everything is enqueued on the stream handed to the call
        cudaStream_t s;
        check_cuda(cudaStreamCreate(&s), "cudaStreamCreate");
        {
            Matrix<double> a(2, 2, {1, 2, 3, 4}, s), b(2, 2, {0, 1, 1, 0}, s);
            EXPECT(gemm(2.0, Transpose::T, a, Transpose::N, b, s).to_host(s) == std::vector<double>({6, 2, 8, 4}), "stream / alpha / T");
        }
        check_cuda(cudaStreamSynchronize(s), "cudaStreamSynchronize");
        cudaStreamDestroy(s);
    } catch (const std::exception &e) {
        printf("FAIL exception: %s\n", e.what());
        return 1;
    }
    if (failures) { printf("hpp_check: %d failure(s)\n", failures); return 1; }
    printf("hpp_check ok (launches %llu)\n", (unsigned long long)accblas_launch_count());
    return 0;
}
