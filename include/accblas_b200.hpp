// accblas_b200.hpp -- C++17 host side above the C ABI (accblas_b200.h), mirroring the reference's operator
// interface for the BLAS hot path: same names, argument order and meaning as
// Data.Array.Accelerate.Numeric.LinearAlgebra (src/Data/Array/Accelerate/Numeric/LinearAlgebra.hs) and its
// BLAS.Level{1,2,3} modules.  The reference's host side is Haskell (compiled code; no GHC in this image), so
// besides the Haskell glue sources (accelerate-blas_b200/haskell/) this header is the compiled-language mirror:
//
//   reference (Haskell)                          here (C++)
//   Numeric e / NumericR (Type.hs:32-51)          numeric_traits<T>, T in float, double, std::complex<float|double>
//   Transpose N | T | H  (Type.hs:78-90)          enum class Transpose { N, T, H }
//   Acc (Matrix e), Acc (Vector e), Acc (Scalar e)  Matrix<T>, Vector<T>, Scalar<T>: dense row-major DEVICE arrays
//   gemm alpha opA a opB b (BLAS/Level3.hs:57)    gemm(alpha, opA, a, opB, b)
//   gemv alpha opA a x     (BLAS/Level2.hs:57)    gemv(alpha, opA, a, x)
//   dotu / dotc            (BLAS/Level1.hs:73,85) dotu(x, y) / dotc(x, y)        (conj on the FIRST argument)
//   sdot / asum / amax / amin (Level1.hs:44-139)  sdot(z, x, y) / asum(x) / amax(x) / amin(x)
//   (<>) (#>) (<#) (<.>) (><) (LinearAlgebra.hs:58-155)   mXm, mXv, vXm, dot, outer  (alpha = 1, as the operators do)
//
// Everything is enqueued on the stream given to the call (default: the legacy default stream) and nothing
// synchronises except the explicit to_host() copies.  Errors are exceptions (accblas::Error carries the status);
// the reference raises Haskell exceptions from its cublas binding.  There is no CPU path behind this header:
// without libaccblas_b200.so the program does not link.
#pragma once
#include <complex>
#include <cstdint>
#include <stdexcept>
#include <string>
#include <utility>
#include <vector>
#include <cuda_runtime_api.h>
#include "accblas_b200.h"

namespace accblas {

enum class Transpose : int { N = ACCBLAS_OP_N, T = ACCBLAS_OP_T, H = ACCBLAS_OP_H };

template <class T> struct numeric_traits;
template <> struct numeric_traits<float> { static constexpr int dtype = ACCBLAS_F32; using real = float; };
template <> struct numeric_traits<double> { static constexpr int dtype = ACCBLAS_F64; using real = double; };
template <> struct numeric_traits<std::complex<float>> { static constexpr int dtype = ACCBLAS_C32; using real = float; };
template <> struct numeric_traits<std::complex<double>> { static constexpr int dtype = ACCBLAS_C64; using real = double; };

class Error : public std::runtime_error {
  public:
    Error(int status, const std::string &where)
        : std::runtime_error(where + ": " + accblas_status_string(status) +
                             (status == ACCBLAS_ERR_CUDA ? std::string(" [") + accblas_last_cuda_error() + "]" : std::string())),
          status_(status) {}
    int status() const { return status_; }
  private:
    int status_;
};
inline void check(int status, const char *where) { if (status != ACCBLAS_OK) throw Error(status, where); }
inline void check_cuda(cudaError_t e, const char *where) {
    if (e != cudaSuccess) throw std::runtime_error(std::string(where) + ": " + cudaGetErrorString(e));
}

// Dense device array of `count` elements, stream-ordered allocation (the role Accelerate's remote memory
// manager plays for the reference, LLVM/PTX/Base.hs:83-120).  Move-only.
template <class T> class DeviceBuffer {
  public:
    DeviceBuffer() = default;
    DeviceBuffer(int64_t count, cudaStream_t s) : count_(count), stream_(s) {
        if (count > 0) check_cuda(cudaMallocAsync(reinterpret_cast<void **>(&ptr_), sizeof(T) * (size_t)count, s), "cudaMallocAsync");
    }
    DeviceBuffer(DeviceBuffer &&o) noexcept { *this = std::move(o); }
    DeviceBuffer &operator=(DeviceBuffer &&o) noexcept {
        if (this != &o) { release(); ptr_ = o.ptr_; count_ = o.count_; stream_ = o.stream_; o.ptr_ = nullptr; o.count_ = 0; }
        return *this;
    }
    DeviceBuffer(const DeviceBuffer &) = delete;
    DeviceBuffer &operator=(const DeviceBuffer &) = delete;
    ~DeviceBuffer() { release(); }
    T *data() { return ptr_; }
    const T *data() const { return ptr_; }
    int64_t count() const { return count_; }
  private:
    void release() { if (ptr_) cudaFreeAsync(ptr_, stream_); ptr_ = nullptr; }
    T *ptr_ = nullptr;
    int64_t count_ = 0;
    cudaStream_t stream_ = nullptr;
};

template <class T> class Vector {
  public:
    Vector() = default;
    explicit Vector(int64_t n, cudaStream_t s = nullptr) : buf_(n, s), n_(n) {}
    Vector(const std::vector<T> &host, cudaStream_t s = nullptr) : Vector((int64_t)host.size(), s) {
        if (n_ > 0) check_cuda(cudaMemcpyAsync(buf_.data(), host.data(), sizeof(T) * (size_t)n_, cudaMemcpyHostToDevice, s), "cudaMemcpyAsync");
    }
    int64_t size() const { return n_; }
    T *data() { return buf_.data(); }
    const T *data() const { return buf_.data(); }
    std::vector<T> to_host(cudaStream_t s = nullptr) const {
        std::vector<T> h((size_t)n_);
        if (n_ > 0) check_cuda(cudaMemcpyAsync(h.data(), buf_.data(), sizeof(T) * (size_t)n_, cudaMemcpyDeviceToHost, s), "cudaMemcpyAsync");
        check_cuda(cudaStreamSynchronize(s), "cudaStreamSynchronize");
        return h;
    }
  private:
    DeviceBuffer<T> buf_;
    int64_t n_ = 0;
};

// Acc (Scalar e): one element in device memory (results of the Level-1 reductions stay on the device, as the
// reference's do; value() is the explicit read-back).
template <class T> class Scalar {
  public:
    explicit Scalar(cudaStream_t s = nullptr) : buf_(1, s) {}
    T *data() { return buf_.data(); }
    const T *data() const { return buf_.data(); }
    T value(cudaStream_t s = nullptr) const {
        T v{};
        check_cuda(cudaMemcpyAsync(&v, buf_.data(), sizeof(T), cudaMemcpyDeviceToHost, s), "cudaMemcpyAsync");
        check_cuda(cudaStreamSynchronize(s), "cudaStreamSynchronize");
        return v;
    }
  private:
    DeviceBuffer<T> buf_;
};

// Dense row-major matrix, tight leading dimension (lda = cols), as every Accelerate array is
// (LLVM/PTX/Level3.hs:64-65).
template <class T> class Matrix {
  public:
    Matrix() = default;
    Matrix(int64_t rows, int64_t cols, cudaStream_t s = nullptr) : buf_(rows * cols, s), rows_(rows), cols_(cols) {}
    Matrix(int64_t rows, int64_t cols, const std::vector<T> &host, cudaStream_t s = nullptr) : Matrix(rows, cols, s) {
        if ((int64_t)host.size() != rows * cols) throw Error(ACCBLAS_ERR_SHAPE, "Matrix(host)");
        if (rows * cols > 0)
            check_cuda(cudaMemcpyAsync(buf_.data(), host.data(), sizeof(T) * host.size(), cudaMemcpyHostToDevice, s), "cudaMemcpyAsync");
    }
    int64_t rows() const { return rows_; }
    int64_t cols() const { return cols_; }
    T *data() { return buf_.data(); }
    const T *data() const { return buf_.data(); }
    std::vector<T> to_host(cudaStream_t s = nullptr) const {
        std::vector<T> h((size_t)(rows_ * cols_));
        if (!h.empty()) check_cuda(cudaMemcpyAsync(h.data(), buf_.data(), sizeof(T) * h.size(), cudaMemcpyDeviceToHost, s), "cudaMemcpyAsync");
        check_cuda(cudaStreamSynchronize(s), "cudaStreamSynchronize");
        return h;
    }
  private:
    DeviceBuffer<T> buf_;
    int64_t rows_ = 0, cols_ = 0;
};

// ------------------------------------------------------------------------------------------------------------
// Level 3 -- gemm alpha opA a opB b = alpha * op(a) * op(b)                       (BLAS/Level3.hs:57-110)
// Stored shapes are pre-op; the result is allocated here, (m x n) with m = rows of op(a), n = columns of op(b)
// (LLVM/PTX/Level3.hs:54-65).  `H` on a real type is `T` (BLAS/Level3.hs:97-100).
// ------------------------------------------------------------------------------------------------------------
template <class T>
Matrix<T> gemm(T alpha, Transpose opA, const Matrix<T> &a, Transpose opB, const Matrix<T> &b, cudaStream_t s = nullptr)
{
    const int64_t m = opA == Transpose::N ? a.rows() : a.cols();
    const int64_t n = opB == Transpose::N ? b.cols() : b.rows();
    Matrix<T> c(m, n, s);
    check(accblas_gemm(s, numeric_traits<T>::dtype, (int)opA, (int)opB, a.rows(), a.cols(), b.rows(), b.cols(), &alpha,
                       a.data(), b.data(), c.data()), "gemm");
    return c;
}

// Level 2 -- gemv alpha opA a x = alpha * op(a) * x                                 (BLAS/Level2.hs:57-93)
template <class T>
Vector<T> gemv(T alpha, Transpose opA, const Matrix<T> &a, const Vector<T> &x, cudaStream_t s = nullptr)
{
    Vector<T> y(opA == Transpose::N ? a.rows() : a.cols(), s);                     // LLVM/PTX/Level2.hs:80-82
    check(accblas_gemv(s, numeric_traits<T>::dtype, (int)opA, a.rows(), a.cols(), &alpha, a.data(), x.data(), x.size(),
                       y.data()), "gemv");
    return y;
}

// Level 1 -- dotu x y = sum x_i y_i ; dotc x y = sum conj(x_i) y_i ; length = min of the two   (BLAS/Level1.hs:73-92)
template <class T> Scalar<T> dotu(const Vector<T> &x, const Vector<T> &y, cudaStream_t s = nullptr)
{
    Scalar<T> d(s);
    check(accblas_dot(s, numeric_traits<T>::dtype, 0, x.size() < y.size() ? x.size() : y.size(), x.data(), y.data(), d.data()), "dotu");
    return d;
}
template <class T> Scalar<T> dotc(const Vector<T> &x, const Vector<T> &y, cudaStream_t s = nullptr)
{
    Scalar<T> d(s);
    check(accblas_dot(s, numeric_traits<T>::dtype, 1, x.size() < y.size() ? x.size() : y.size(), x.data(), y.data(), d.data()), "dotc");
    return d;
}
// compensated dotu / dotc: exact product splitting + the reference's KBN combine (Numeric/Sum.hs:119-125)
template <class T> Scalar<T> dot_compensated(const Vector<T> &x, const Vector<T> &y, bool conj = false, cudaStream_t s = nullptr)
{
    Scalar<T> d(s);
    check(accblas_dot_compensated(s, numeric_traits<T>::dtype, conj ? 1 : 0, x.size() < y.size() ? x.size() : y.size(), x.data(),
                                  y.data(), d.data()), "dot_compensated");
    return d;
}
// sdot z x y = z + sum x_i y_i accumulated in double precision                      (BLAS/Level1.hs:44-62)
template <class T> Scalar<T> sdot(T z, const Vector<T> &x, const Vector<T> &y, cudaStream_t s = nullptr)
{
    Scalar<T> d(s);
    check(accblas_sdot(s, numeric_traits<T>::dtype, x.size() < y.size() ? x.size() : y.size(), &z, x.data(), y.data(), d.data()), "sdot");
    return d;
}
// asum x = sum |re x_i| + |im x_i| (a real scalar)                                   (BLAS/Level1.hs:100-110)
template <class T> Scalar<typename numeric_traits<T>::real> asum(const Vector<T> &x, cudaStream_t s = nullptr)
{
    Scalar<typename numeric_traits<T>::real> d(s);
    check(accblas_asum(s, numeric_traits<T>::dtype, x.size(), x.data(), d.data()), "asum");
    return d;
}
// amax / amin: index of the element of largest / smallest magnitude (last one among ties)   (BLAS/Level1.hs:118-141)
template <class T> Scalar<int64_t> amax(const Vector<T> &x, cudaStream_t s = nullptr)
{
    Scalar<int64_t> d(s);
    check(accblas_amax(s, numeric_traits<T>::dtype, x.size(), x.data(), d.data()), "amax");
    return d;
}
template <class T> Scalar<int64_t> amin(const Vector<T> &x, cudaStream_t s = nullptr)
{
    Scalar<int64_t> d(s);
    check(accblas_amin(s, numeric_traits<T>::dtype, x.size(), x.data(), d.data()), "amin");
    return d;
}

// ------------------------------------------------------------------------------------------------------------
// The operators of Data.Array.Accelerate.Numeric.LinearAlgebra (alpha = 1, op = N / T as the operators fix them)
// ------------------------------------------------------------------------------------------------------------
template <class T> Matrix<T> mXm(const Matrix<T> &a, const Matrix<T> &b, cudaStream_t s = nullptr)     // (<>)  :153-155
{ return gemm(T(1), Transpose::N, a, Transpose::N, b, s); }
template <class T> Vector<T> mXv(const Matrix<T> &a, const Vector<T> &x, cudaStream_t s = nullptr)     // (#>)  :99-101
{ return gemv(T(1), Transpose::N, a, x, s); }
template <class T> Vector<T> vXm(const Vector<T> &x, const Matrix<T> &a, cudaStream_t s = nullptr)     // (<#)  :119-121
{ return gemv(T(1), Transpose::T, a, x, s); }
template <class T> Scalar<T> dot(const Vector<T> &x, const Vector<T> &y, cudaStream_t s = nullptr)     // (<.>) :58-60
{ return dotu(x, y, s); }
// (><) :73-78 -- outer product: the reference reshapes x to (n x 1), y to (1 x m) and multiplies; here the
// k = 1 product is served by the rank-1 store-stream kernel
template <class T> Matrix<T> outer(const Vector<T> &x, const Vector<T> &y, cudaStream_t s = nullptr)
{
    Matrix<T> c(x.size(), y.size(), s);
    const T one(1);
    check(accblas_gemm(s, numeric_traits<T>::dtype, ACCBLAS_OP_N, ACCBLAS_OP_N, x.size(), 1, 1, y.size(), &one, x.data(),
                       y.data(), c.data()), "outer");
    return c;
}

}  // namespace accblas
