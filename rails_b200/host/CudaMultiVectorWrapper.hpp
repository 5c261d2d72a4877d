// MultiVector concept of RAILS::Solver for the B200 backend.
//
// Same member functions, argument meaning and error behaviour as the
// reference implementations (StlWrapper, /root/reference/src/StlWrapper.hpp:11-90,
// and Epetra_MultiVectorWrapper, src/Epetra_MultiVectorWrapper.hpp), so that
// Solver<CudaOperatorWrapper, CudaMultiVectorWrapper, HostDenseMatrix>
// instantiates from the unmodified src/LyapunovSolver.hpp.  The data lives in
// HBM behind a rails_mv_t (include/rails_cuda.h); this class is host-side
// bookkeeping only -- sizes, views, ownership -- and every arithmetic member
// is one or two calls into the C ABI.
//
// Ownership/aliasing rules kept from the reference:
//   * copy constructor is DEEP, operator= is SHALLOW unless the target is a
//     view, in which case the columns are copied in place
//     (src/StlWrapper.cpp:31-44, 65-121);
//   * view(a, b) shares the parent's storage (here through a shared_ptr, so a
//     view can never dangle);
//   * resize(n) within capacity only edits the logical size (:219-263);
//   * results are returned by value; a real move constructor/assignment makes
//     `MultiVector X = std::move(expr)` free.
//
// Lazy products: `V * C`, `B * Y` and `alpha * X` return an object that only
// records the operation.  `y += V * C`, `y -= alpha * x` and `view = V * C`
// then run as ONE fused kernel with no m-sized temporary; any other use
// materialises the product first.  The fused kernels add the finished sum to
// y in a separately rounded step, so the bits are those of the unfused
// sequence.
#ifndef RAILS_B200_CUDAMULTIVECTORWRAPPER_HPP
#define RAILS_B200_CUDAMULTIVECTORWRAPPER_HPP

#include "HostDenseMatrix.hpp"
#include "rails_cuda.h"

#include <memory>
#include <vector>

namespace RAILS
{

class CudaOperatorWrapper;

class CudaMultiVectorWrapper
{
public:
    struct Lazy;
    struct Buffer
    {
        rails_mv_t h;
        // products that still READ this buffer; they are evaluated before the
        // buffer is written to, so laziness is never observable
        std::vector<std::weak_ptr<Lazy> > readers;
        explicit Buffer(rails_mv_t handle) : h(handle) {}
        ~Buffer() { rails_mv_free(h); }
        Buffer(Buffer const &) = delete;
        Buffer &operator=(Buffer const &) = delete;
        void before_write(int col0, int ncols);
    };

    // test-only element access (the reference keeps operator()(i,j) protected
    // in the Epetra wrapper and re-exposes it for tests,
    // test/Epetra_TestableWrappers.hpp:62-70)
    class Element
    {
        rails_mv_t h_;
        long long row_;
        int col_;
    public:
        Element(rails_mv_t h, long long row, int col) : h_(h), row_(row), col_(col) {}
        operator double() const;
        Element &operator=(double v);
        Element &operator=(Element const &o) { return *this = (double)o; }
        Element &operator*=(double v) { return *this = (double)*this * v; }
        Element &operator+=(double v) { return *this = (double)*this + v; }
        Element &operator-=(double v) { return *this = (double)*this - v; }
    };

    // pending product  alpha * src[:, src_col0 : src_col0+src_k) * coef
    struct Lazy
    {
        std::shared_ptr<Buffer> src;
        int src_col0 = 0, src_k = 0;
        bool has_coef = false;                 // false: plain alpha * src
        HostDenseMatrix coef;                  // host coefficients (k x c), or
        std::shared_ptr<Buffer> dev_coef;      // replicated device block
        int dev_coef_col0 = 0;
        double alpha = 1.0;
        int m = 0, n = 0;                      // shape of the result
        bool replicated = false;
        std::shared_ptr<Buffer> result;        // set once evaluated
        void evaluate();
        bool reads(Buffer const *b, int col0, int ncols) const;
    };

private:
    mutable std::shared_ptr<Buffer> buf_;
    int col0_;            // first column of this object inside buf_
    int m_;               // GLOBAL number of rows
    int n_;               // logical number of columns
    int n_max_;           // columns available from col0_
    int orthogonalized_;  // leading columns known to be orthonormal
    bool is_view_;
    bool transpose_;
    bool replicated_;     // identical on all ranks (Epetra_LocalMap analogue)

    mutable std::shared_ptr<Lazy> lazy_;

    void allocate(int m, int n, bool replicated);
    void materialize() const;
    void assign_columns(CudaMultiVectorWrapper const &other);
    void add_scaled(CudaMultiVectorWrapper const &other, double alpha);
    void will_write();
    static CudaMultiVectorWrapper make_lazy(std::shared_ptr<Lazy> const &l);
    rails_mv_t handle() const;

    friend class CudaOperatorWrapper;
    friend CudaMultiVectorWrapper operator*(double d, CudaMultiVectorWrapper const &other);

public:
    CudaMultiVectorWrapper();
    CudaMultiVectorWrapper(CudaMultiVectorWrapper const &other);
    CudaMultiVectorWrapper(CudaMultiVectorWrapper &&other) noexcept;
    CudaMultiVectorWrapper(CudaMultiVectorWrapper const &other, int n);
    CudaMultiVectorWrapper(int m, int n);
    static CudaMultiVectorWrapper replicated(int m, int n);
    // m x n block from host memory (column-major, leading dimension ldh >= m;
    // this rank's rows are taken from the global block)
    static CudaMultiVectorWrapper from_host(int m, int n, const double *host, long long ldh);

    virtual ~CudaMultiVectorWrapper() {}

    CudaMultiVectorWrapper &operator=(CudaMultiVectorWrapper &other);
    CudaMultiVectorWrapper &operator=(CudaMultiVectorWrapper const &other);
    CudaMultiVectorWrapper &operator=(CudaMultiVectorWrapper &&other);

    CudaMultiVectorWrapper &operator=(double other);

    CudaMultiVectorWrapper &operator*=(double other);
    CudaMultiVectorWrapper &operator/=(double other);

    CudaMultiVectorWrapper &operator-=(CudaMultiVectorWrapper const &other);
    CudaMultiVectorWrapper &operator+=(CudaMultiVectorWrapper const &other);

    CudaMultiVectorWrapper operator+(CudaMultiVectorWrapper const &other) const;
    CudaMultiVectorWrapper operator*(CudaMultiVectorWrapper const &other) const;
    CudaMultiVectorWrapper operator*(HostDenseMatrix const &other) const;

    Element operator()(int i, int j = 0);
    double operator()(int i, int j = 0) const;

    void resize(int n);
    void resize(int m, int n);

    double norm() const;
    double norm_inf() const;

    void orthogonalize();

    CudaMultiVectorWrapper view(int m = -1, int n = -1);
    CudaMultiVectorWrapper view(int m = -1, int n = -1) const;
    CudaMultiVectorWrapper copy() const;

    void push_back(CudaMultiVectorWrapper const &other);

    int M() const;
    int N() const;

    HostDenseMatrix dot(CudaMultiVectorWrapper const &other) const;

    void random();

    CudaMultiVectorWrapper transpose() const;

    // copy this rank's rows to host memory (column-major, ldh >= local rows)
    int to_host(double *host, long long ldh) const;
    long long local_rows() const;
    bool is_replicated() const { return replicated_; }
    // device handle + first column, for the fused paths of the solver loop
    rails_mv_t device_handle() const { return handle(); }
    int device_col0() const { materialize(); return col0_; }
};

CudaMultiVectorWrapper operator*(double d, CudaMultiVectorWrapper const &other);

}

#endif
