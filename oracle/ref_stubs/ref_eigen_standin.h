// ref_eigen_standin.h -- TEST INFRASTRUCTURE.  Minimal stand-in for the parts of Eigen that the reference's hot-path
// translation units use, so that those files compile UNMODIFIED from /root/reference (Eigen is not installed here and
// is not part of the reference tree).  Written from the call sites, C++03 like the reference:
//   Vector3i/3f, Matrix3f, VectorXf, MatrixXf            grid.h:23-25, dfields.cpp:254-309, mcubes.cpp:9-152
//   SelfAdjointEigenSolver<Matrix3f>                      dfields.cpp:293-309
//   JacobiSVD<MatrixXf> (+ setThreshold, solve)           mcubes.cpp:142-152
//   Triplet, SparseMatrix<float>, SparseVector<float>     primeqp.cpp:101-110,190-210, quadprog.cpp:8-42
// What this does NOT pin (named in DESIGN.md): the floating-point path inside Eigen's eigen-solver and SVD, and the
// fact that Eigen's sparse product / setFromTriplets return row indexes ascending inside each column (the stand-in
// sorts them; every value on that path is a small exact integer, so only the ORDER is an assumption).
//   default build: the eigen-solver and the least-squares solve use the arithmetic the product defines in
//                  include/mrb_eig3.h, so reference code + stand-in can be compared with the GPU path bit for bit;
//   -DREF_STANDIN_DOUBLE: an independent double-precision Jacobi eigen-solver / one-sided Jacobi SVD rounded to fp32
//                  (what an exact solver returns), used to bound how far the mrb_eig3.h arithmetic can be from Eigen's.
#ifndef REF_EIGEN_STANDIN_H
#define REF_EIGEN_STANDIN_H

#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <fstream>
#include <iostream>
#include <string>
#include <vector>
#include <math.h>

#ifndef REF_STANDIN_DOUBLE
#include "mrb_eig3.h"
#endif

namespace Eigen {

const int Dynamic = -1;
enum { ComputeFullU = 4, ComputeThinU = 8, ComputeFullV = 16, ComputeThinV = 32 };

namespace standin {
template <typename T, int N> struct Store {
    T a[N];
    Store() { for (int i = 0; i < N; i++) a[i] = T(); }
    T* p() { return a; }
    const T* p() const { return a; }
    void alloc(int) {}
};
template <typename T> struct Store<T, -1> {
    std::vector<T> v;
    T* p() { return v.empty() ? 0 : &v[0]; }
    const T* p() const { return v.empty() ? 0 : &v[0]; }
    void alloc(int n) { v.assign(n, T()); }
};
}  // namespace standin

template <typename T, int R, int C> class Matrix {
    standin::Store<T, (R == Dynamic || C == Dynamic) ? -1 : R * C> s_;
    int r_, c_;

public:
    Matrix() : r_(R == Dynamic ? 0 : R), c_(C == Dynamic ? 0 : C) {}
    explicit Matrix(int n) : r_(R == Dynamic ? n : R), c_(C == Dynamic ? (R == Dynamic ? 1 : n) : C) { s_.alloc(r_ * c_); }
    Matrix(int r, int c) : r_(r), c_(c) { s_.alloc(r * c); }
    template <int R2, int C2> Matrix(const Matrix<T, R2, C2>& o) : r_(o.rows()), c_(o.cols()) {
        s_.alloc(r_ * c_);
        for (int j = 0; j < c_; j++) for (int i = 0; i < r_; i++) (*this)(i, j) = o(i, j);
    }
    int rows() const { return r_; }
    int cols() const { return c_; }
    int size() const { return r_ * c_; }
    void resize(int n) { if (C == 1 || C == Dynamic) { r_ = n; c_ = 1; } else { r_ = 1; c_ = n; } s_.alloc(n); }
    void resize(int r, int c) { r_ = r; c_ = c; s_.alloc(r * c); }
    // column-major like Eigen (never observable through the reference's call sites)
    T& operator()(int i, int j) { return s_.p()[j * r_ + i]; }
    const T& operator()(int i, int j) const { return s_.p()[j * r_ + i]; }
    T& operator()(int i) { return s_.p()[i]; }
    const T& operator()(int i) const { return s_.p()[i]; }
    T& operator[](int i) { return s_.p()[i]; }
    const T& operator[](int i) const { return s_.p()[i]; }
    Matrix<T, R, 1> col(int j) const {
        Matrix<T, R, 1> o(r_, 1);
        for (int i = 0; i < r_; i++) o(i, 0) = (*this)(i, j);
        return o;
    }
    // "m << a, b, c, ..." fills row by row (dfields.cpp:272)
    class CommaInit {
        Matrix* m_; int k_;
    public:
        CommaInit(Matrix* m, T v) : m_(m), k_(0) { put(v); }
        void put(T v) { (*m_)(k_ / m_->cols(), k_ % m_->cols()) = v; k_++; }
        CommaInit& operator,(T v) { put(v); return *this; }
    };
    CommaInit operator<<(T v) { return CommaInit(this, v); }
};

typedef Matrix<int, 3, 1> Vector3i;
typedef Matrix<float, 3, 1> Vector3f;
typedef Matrix<float, 3, 3> Matrix3f;
typedef Matrix<float, Dynamic, 1> VectorXf;
typedef Matrix<float, Dynamic, Dynamic> MatrixXf;

#ifdef REF_STANDIN_DOUBLE
namespace standin {
// cyclic Jacobi in double until the off-diagonal mass vanishes; eigenvectors in the columns of v
inline void jacobi3d(double a[3][3], double w[3], double v[3][3]) {
    for (int i = 0; i < 3; i++) for (int j = 0; j < 3; j++) v[i][j] = (i == j);
    for (int sweep = 0; sweep < 60; sweep++) {
        double off = fabs(a[0][1]) + fabs(a[0][2]) + fabs(a[1][2]);
        if (off < 1e-300) break;
        for (int p = 0; p < 2; p++) for (int q = p + 1; q < 3; q++) {
            if (a[p][q] == 0.0) continue;
            double theta = (a[q][q] - a[p][p]) / (2.0 * a[p][q]);
            double t = (theta >= 0 ? 1.0 : -1.0) / (fabs(theta) + sqrt(theta * theta + 1.0));
            double c = 1.0 / sqrt(t * t + 1.0), s = t * c;
            for (int k = 0; k < 3; k++) { double akp = a[k][p], akq = a[k][q]; a[k][p] = c * akp - s * akq; a[k][q] = s * akp + c * akq; }
            for (int k = 0; k < 3; k++) { double apk = a[p][k], aqk = a[q][k]; a[p][k] = c * apk - s * aqk; a[q][k] = s * apk + c * aqk; }
            for (int k = 0; k < 3; k++) { double vkp = v[k][p], vkq = v[k][q]; v[k][p] = c * vkp - s * vkq; v[k][q] = s * vkp + c * vkq; }
        }
    }
    for (int i = 0; i < 3; i++) w[i] = a[i][i];
}
}  // namespace standin
#endif

// SelfAdjointEigenSolver: eigenvalues ascending, eigenvectors in the matching columns (Eigen's documented contract).
template <typename M> class SelfAdjointEigenSolver {
    Vector3f w_;
    Matrix3f v_;

public:
    SelfAdjointEigenSolver() {}
    SelfAdjointEigenSolver& compute(const Matrix3f& m) {
        int order[3] = {0, 1, 2};
#ifndef REF_STANDIN_DOUBLE
        float a[9], w[3], v[9];
        for (int r = 0; r < 3; r++) for (int c = 0; c < 3; c++) a[r * 3 + c] = m(r, c);
        mrb_eig3_jacobi(a, w, v);
#else
        double a[3][3], w[3], vv[3][3], v[9];
        for (int r = 0; r < 3; r++) for (int c = 0; c < 3; c++) a[r][c] = m(r, c);
        standin::jacobi3d(a, w, vv);
        for (int r = 0; r < 3; r++) for (int c = 0; c < 3; c++) v[r * 3 + c] = vv[r][c];
#endif
        for (int i = 1; i < 3; i++)
            for (int j = i; j > 0 && w[order[j]] < w[order[j - 1]]; j--) std::swap(order[j], order[j - 1]);
        for (int i = 0; i < 3; i++) {
            w_[i] = (float)w[order[i]];
            for (int r = 0; r < 3; r++) v_(r, i) = (float)v[r * 3 + order[i]];
        }
        return *this;
    }
    const Vector3f& eigenvalues() const { return w_; }
    const Matrix3f& eigenvectors() const { return v_; }
};

// JacobiSVD of an n x 3 matrix; only singularValues(), setThreshold() and solve() are used (mcubes.cpp:144-149).
// Eigen's rank rule: singular values > max(sigma_0 * threshold, tiny) count; default threshold = epsilon * min(rows, cols).
template <typename M> class JacobiSVD {
    int n_;
    std::vector<float> m_;    // row-major n x 3
    VectorXf sv_;             // descending
    bool has_thr_;
    float thr_;
#ifndef REF_STANDIN_DOUBLE
    float w_[3], v_[9], sg_[3];
#else
    double sd_[3], vd_[3][3];
    std::vector<double> u_;   // n x 3, column j = left singular vector of sd_[j]
#endif
public:
    JacobiSVD(const MatrixXf& m, unsigned int = 0) : n_(m.rows()), has_thr_(false), thr_(0.0f) {
        if (m.cols() != 3) { fprintf(stderr, "JacobiSVD stand-in: only n x 3 matrices\n"); exit(1); }
        m_.resize(n_ * 3);
        for (int i = 0; i < n_; i++) for (int c = 0; c < 3; c++) m_[i * 3 + c] = m(i, c);
        int nsv = n_ < 3 ? n_ : 3;
        sv_.resize(nsv);
#ifndef REF_STANDIN_DOUBLE
        float a[9] = {0, 0, 0, 0, 0, 0, 0, 0, 0};
        for (int i = 0; i < n_; i++)
            for (int r = 0; r < 3; r++)
                for (int c = 0; c < 3; c++) a[r * 3 + c] = a[r * 3 + c] + m_[i * 3 + r] * m_[i * 3 + c];
        mrb_eig3_jacobi(a, w_, v_);
        float s[3];
        for (int i = 0; i < 3; i++) { sg_[i] = sqrtf(w_[i] > 0.0f ? w_[i] : 0.0f); s[i] = sg_[i]; }
        std::sort(s, s + 3);
        for (int i = 0; i < nsv; i++) sv_[i] = s[2 - i];
#else
        // one-sided (Hestenes) Jacobi on the columns of a double copy
        std::vector<double> a(n_ * 3);
        for (int i = 0; i < n_ * 3; i++) a[i] = m_[i];
        for (int i = 0; i < 3; i++) for (int j = 0; j < 3; j++) vd_[i][j] = (i == j);
        for (int sweep = 0; sweep < 60; sweep++) {
            double off = 0;
            for (int p = 0; p < 2; p++) for (int q = p + 1; q < 3; q++) {
                double al = 0, be = 0, ga = 0;
                for (int i = 0; i < n_; i++) { al += a[i * 3 + p] * a[i * 3 + p]; be += a[i * 3 + q] * a[i * 3 + q]; ga += a[i * 3 + p] * a[i * 3 + q]; }
                off += fabs(ga);
                if (ga == 0.0) continue;
                double zeta = (be - al) / (2.0 * ga);
                double t = (zeta >= 0 ? 1.0 : -1.0) / (fabs(zeta) + sqrt(1.0 + zeta * zeta));
                double c = 1.0 / sqrt(1.0 + t * t), s = c * t;
                for (int i = 0; i < n_; i++) { double x = a[i * 3 + p], y = a[i * 3 + q]; a[i * 3 + p] = c * x - s * y; a[i * 3 + q] = s * x + c * y; }
                for (int i = 0; i < 3; i++) { double x = vd_[i][p], y = vd_[i][q]; vd_[i][p] = c * x - s * y; vd_[i][q] = s * x + c * y; }
            }
            if (off < 1e-300) break;
        }
        u_.assign(n_ * 3, 0.0);
        for (int j = 0; j < 3; j++) {
            double nn = 0;
            for (int i = 0; i < n_; i++) nn += a[i * 3 + j] * a[i * 3 + j];
            sd_[j] = sqrt(nn);
            for (int i = 0; i < n_; i++) u_[i * 3 + j] = sd_[j] > 0 ? a[i * 3 + j] / sd_[j] : 0.0;
        }
        double s[3] = {sd_[0], sd_[1], sd_[2]};
        std::sort(s, s + 3);
        for (int i = 0; i < nsv; i++) sv_[i] = (float)s[2 - i];
#endif
    }
    const VectorXf& singularValues() const { return sv_; }
    JacobiSVD& setThreshold(float t) { has_thr_ = true; thr_ = t; return *this; }
    Vector3f solve(const VectorXf& vec) const {
        Vector3f out;
        const float thr = has_thr_ ? thr_ : 3.0f * 1.1920929e-7f;
#ifndef REF_STANDIN_DOUBLE
        float b[3] = {0, 0, 0}, p[3] = {0, 0, 0};
        for (int i = 0; i < n_; i++) for (int r = 0; r < 3; r++) b[r] = b[r] + m_[i * 3 + r] * vec[i];
        float smax = sg_[0];
        for (int i = 1; i < 3; i++) if (sg_[i] > smax) smax = sg_[i];
        for (int i = 0; i < 3; i++) {
            if (sg_[i] > thr * smax) {
                float proj = v_[0 * 3 + i] * b[0] + v_[1 * 3 + i] * b[1] + v_[2 * 3 + i] * b[2];
                float coef = proj / w_[i];
                p[0] = p[0] + v_[0 * 3 + i] * coef; p[1] = p[1] + v_[1 * 3 + i] * coef; p[2] = p[2] + v_[2 * 3 + i] * coef;
            }
        }
        out[0] = p[0]; out[1] = p[1]; out[2] = p[2];
#else
        double smax = std::max(sd_[0], std::max(sd_[1], sd_[2])), p[3] = {0, 0, 0};
        for (int j = 0; j < 3; j++) {
            if ((float)sd_[j] > thr * (float)smax) {
                double proj = 0;
                for (int i = 0; i < n_; i++) proj += u_[i * 3 + j] * vec[i];
                for (int r = 0; r < 3; r++) p[r] += vd_[r][j] * proj / sd_[j];
            }
        }
        out[0] = (float)p[0]; out[1] = (float)p[1]; out[2] = (float)p[2];
#endif
        return out;
    }
};

// ---- sparse (exact small-integer arithmetic on this path; compressed column storage, rows ascending) ----
template <typename T> class Triplet {
    int r_, c_; T v_;
public:
    Triplet() : r_(0), c_(0), v_(T()) {}
    Triplet(int r, int c, const T& v = T()) : r_(r), c_(c), v_(v) {}
    int row() const { return r_; }
    int col() const { return c_; }
    const T& value() const { return v_; }
};

template <typename T> class SparseVector {
    int n_;
    std::vector<int> idx_;
    std::vector<T> val_;
public:
    SparseVector() : n_(0) {}
    explicit SparseVector(int n) : n_(n) {}
    void push(int i, T v) { idx_.push_back(i); val_.push_back(v); }
    int size() const { return n_; }
    class InnerIterator {
        const SparseVector* s_; size_t k_;
    public:
        InnerIterator(const SparseVector& s, int = 0) : s_(&s), k_(0) {}
        operator bool() const { return k_ < s_->idx_.size(); }
        InnerIterator& operator++() { k_++; return *this; }
        int index() const { return s_->idx_[k_]; }
        T value() const { return s_->val_[k_]; }
    };
};

template <typename T> class SparseMatrix {
    int rows_, cols_;
    std::vector<int> outer_;   // cols_ + 1
    std::vector<int> inner_;
    std::vector<T> val_;

    struct DenseDiag {
        std::vector<T> d;
        SparseVector<T> sparseView() const {        // keeps entries that are not exactly 0
            SparseVector<T> s((int)d.size());
            for (size_t i = 0; i < d.size(); i++) if (d[i] != T(0)) s.push((int)i, d[i]);
            return s;
        }
    };
    struct Entry { int r; T v; bool operator<(const Entry& o) const { return r < o.r; } };

public:
    SparseMatrix() : rows_(0), cols_(0), outer_(1, 0) {}
    SparseMatrix(int r, int c) : rows_(r), cols_(c), outer_(c + 1, 0) {}
    int rows() const { return rows_; }
    int cols() const { return cols_; }
    int outerSize() const { return cols_; }
    int nonZeros() const { return (int)inner_.size(); }

    template <typename It> void setFromTriplets(It b, It e) {     // duplicates are summed
        std::vector<std::vector<Entry> > colv(cols_);
        for (It it = b; it != e; ++it) {
            if (it->row() < 0 || it->row() >= rows_ || it->col() < 0 || it->col() >= cols_) {
                fprintf(stderr, "SparseMatrix stand-in: triplet (%d,%d) outside %d x %d\n", it->row(), it->col(), rows_, cols_);
                exit(1);
            }
            Entry en; en.r = it->row(); en.v = it->value();
            colv[it->col()].push_back(en);
        }
        build(colv);
    }
    SparseMatrix transpose() const {
        SparseMatrix t(cols_, rows_);
        std::vector<std::vector<Entry> > colv(rows_);
        for (int c = 0; c < cols_; c++)
            for (int k = outer_[c]; k < outer_[c + 1]; k++) { Entry en; en.r = c; en.v = val_[k]; colv[inner_[k]].push_back(en); }
        t.build(colv);
        return t;
    }
    friend SparseMatrix operator*(const SparseMatrix& a, const SparseMatrix& b) {
        SparseMatrix o(a.rows_, b.cols_);
        std::vector<std::vector<Entry> > colv(b.cols_);
        for (int c = 0; c < b.cols_; c++)
            for (int k = b.outer_[c]; k < b.outer_[c + 1]; k++) {
                const int j = b.inner_[k];
                const T bv = b.val_[k];
                for (int l = a.outer_[j]; l < a.outer_[j + 1]; l++) { Entry en; en.r = a.inner_[l]; en.v = a.val_[l] * bv; colv[c].push_back(en); }
            }
        o.build(colv);
        return o;
    }
    friend SparseMatrix operator-(const SparseMatrix& a, const SparseMatrix& b) {
        SparseMatrix o(a.rows_, a.cols_);
        std::vector<std::vector<Entry> > colv(a.cols_);
        for (int c = 0; c < a.cols_; c++) {
            for (int k = a.outer_[c]; k < a.outer_[c + 1]; k++) { Entry en; en.r = a.inner_[k]; en.v = a.val_[k]; colv[c].push_back(en); }
            for (int k = b.outer_[c]; k < b.outer_[c + 1]; k++) { Entry en; en.r = b.inner_[k]; en.v = -b.val_[k]; colv[c].push_back(en); }
        }
        o.build(colv);
        return o;
    }
    void prune(const T& reference, const T& epsilon) {              // drops |v| <= |reference| * epsilon
        std::vector<int> no(cols_ + 1, 0), ni; std::vector<T> nv;
        const T lim = (reference < 0 ? -reference : reference) * epsilon;
        for (int c = 0; c < cols_; c++) {
            for (int k = outer_[c]; k < outer_[c + 1]; k++) {
                T av = val_[k] < 0 ? -val_[k] : val_[k];
                if (av > lim) { ni.push_back(inner_[k]); nv.push_back(val_[k]); }
            }
            no[c + 1] = (int)ni.size();
        }
        outer_.swap(no); inner_.swap(ni); val_.swap(nv);
    }
    DenseDiag diagonal() const {
        DenseDiag dd;
        dd.d.assign(rows_ < cols_ ? rows_ : cols_, T(0));
        for (int c = 0; c < (int)dd.d.size(); c++)
            for (int k = outer_[c]; k < outer_[c + 1]; k++) if (inner_[k] == c) dd.d[c] = val_[k];
        return dd;
    }
    class InnerIterator {
        const SparseMatrix* m_; int k_, end_, outer_;
    public:
        InnerIterator(const SparseMatrix& m, int outer) : m_(&m), k_(m.outer_[outer]), end_(m.outer_[outer + 1]), outer_(outer) {}
        operator bool() const { return k_ < end_; }
        InnerIterator& operator++() { k_++; return *this; }
        int row() const { return m_->inner_[k_]; }
        int col() const { return outer_; }
        int index() const { return m_->inner_[k_]; }
        T value() const { return m_->val_[k_]; }
    };

private:
    void build(std::vector<std::vector<Entry> >& colv) {           // sort rows, sum duplicates (structural zeros are kept)
        outer_.assign(cols_ + 1, 0); inner_.clear(); val_.clear();
        for (int c = 0; c < cols_; c++) {
            std::vector<Entry>& v = colv[c];
            std::stable_sort(v.begin(), v.end());
            for (size_t k = 0; k < v.size();) {
                T s = v[k].v; size_t l = k + 1;
                while (l < v.size() && v[l].r == v[k].r) { s = s + v[l].v; l++; }
                inner_.push_back(v[k].r); val_.push_back(s);
                k = l;
            }
            outer_[c + 1] = (int)inner_.size();
        }
    }
};

}  // namespace Eigen
#endif
