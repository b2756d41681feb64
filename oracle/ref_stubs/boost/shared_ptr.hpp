// TEST INFRASTRUCTURE: stand-in for boost::shared_ptr (C++03, reference-counted), enough for the reference's
// gridPtr / bandsPtr / qp_argsPtr / SparseMatrixPtr typedefs (grid.h:86, narrowBand.h:15, primeqp.h:17,27).
#ifndef REF_STANDIN_BOOST_SHARED_PTR
#define REF_STANDIN_BOOST_SHARED_PTR
namespace boost {
template <typename T> class shared_ptr {
    T* p_;
    long* n_;
    void drop() { if (n_ && --*n_ == 0) { delete p_; delete n_; } p_ = 0; n_ = 0; }
public:
    shared_ptr() : p_(0), n_(0) {}
    explicit shared_ptr(T* p) : p_(p), n_(p ? new long(1) : 0) {}
    shared_ptr(const shared_ptr& o) : p_(o.p_), n_(o.n_) { if (n_) ++*n_; }
    shared_ptr& operator=(const shared_ptr& o) {
        if (this != &o) { drop(); p_ = o.p_; n_ = o.n_; if (n_) ++*n_; }
        return *this;
    }
    ~shared_ptr() { drop(); }
    T* operator->() const { return p_; }
    T& operator*() const { return *p_; }
    T* get() const { return p_; }
    operator bool() const { return p_ != 0; }
};
}  // namespace boost
#endif
