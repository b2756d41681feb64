// mesh_reconstruction_compat.h -- the reference's C++ module API (same names, argument meaning and
// error behaviour) on top of the C-ABI of libmrb.so.  Header-only; link with -lmrb.
//
// What a user of CRLab/Mesh_Reconstruction includes instead of grid.h / dfields.h / narrowBand.h /
// primeqp.h / quadprog.h / mcubes.h / hasher.h (thin forwarding headers with those names sit next
// to this file).  No PCL, Boost or Eigen is needed: boost::shared_ptr -> std::shared_ptr and the
// two Eigen vector types the signatures mention are replaced by minimal stand-ins in namespace Eigen.
//
// Error behaviour follows the reference: message on cerr, then exit(1)
// (grid.cpp:20-23, dfields.cu:7-16, mcubes.cpp:652-656).
#pragma once

#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <iostream>
#include <memory>
#include <vector>

#include "mrb.h"

namespace Eigen {   // stand-ins for the two Eigen types that appear in the reference signatures
struct Vector3i {
    int v[3] = {0, 0, 0};
    Vector3i() {}
    Vector3i(int a, int b, int c) { v[0] = a; v[1] = b; v[2] = c; }
    int& operator[](int i) { return v[i]; }
    const int& operator[](int i) const { return v[i]; }
};
struct Vector3f {
    float v[3] = {0, 0, 0};
    Vector3f() {}
    Vector3f(float a, float b, float c) { v[0] = a; v[1] = b; v[2] = c; }
    float& operator[](int i) { return v[i]; }
    const float& operator[](int i) const { return v[i]; }
};
}  // namespace Eigen

namespace mrbcompat {

inline mrb_ctx* ctx() {      // one context per thread, created on first use (device 0 or $MRB_DEVICE)
    static thread_local mrb_ctx* c = nullptr;
    if (!c) {
        const char* e = std::getenv("MRB_DEVICE");
        if (mrb_ctx_create(e ? std::atoi(e) : 0, &c) != MRB_OK) {
            std::cerr << mrb_last_error() << std::endl;
            std::exit(EXIT_FAILURE);
        }
    }
    return c;
}
inline void check(int rc) {
    if (rc != MRB_OK) {
        std::cerr << mrb_last_error() << std::endl;
        std::exit(EXIT_FAILURE);
    }
}

}  // namespace mrbcompat

// ---- grid (grid.h:16-104) ------------------------------------------------------------------------------
class grid {
  private:
    std::vector<float> voxels;      // [x][y][z], z contiguous (grid.cpp:41-48)

  public:
    Eigen::Vector3i dims;
    Eigen::Vector3f shift;
    Eigen::Vector3f scale;
    int pad;

    class y_z {
        float* p; int dy, dz;
      public:
        class z {
            float* p; int dz;
          public:
            z(float* p_, int dz_) : p(p_), dz(dz_) {}
            float& operator[](int index) {
                if (index < 0 || index >= dz) { std::cerr << "z index of grid out of range [0:" << dz << "): " << index << std::endl; std::exit(1); }
                return p[index];
            }
        };
        y_z(float* p_, int dy_, int dz_) : p(p_), dy(dy_), dz(dz_) {}
        z operator[](int index) {
            if (index < 0 || index >= dy) { std::cerr << "y index of grid out of range [0:" << dy << "): " << index << std::endl; std::exit(1); }
            return z(p + (size_t)index * dz, dz);
        }
    };
    y_z operator[](int index) {
        if (index < 0 || index >= dims[0]) { std::cerr << "x index of grid out of range [0:" << dims[0] << "): " << index << std::endl; std::exit(1); }
        return y_z(voxels.data() + (size_t)index * dims[1] * dims[2], dims[1], dims[2]);
    }

    grid(const Eigen::Vector3i& dims_in, const Eigen::Vector3f& scale_in, const Eigen::Vector3f& shift_in, const int pad_in)
        : voxels((size_t)dims_in[0] * dims_in[1] * dims_in[2], 0.0f), dims(dims_in), shift(shift_in), scale(scale_in), pad(pad_in) {}
    grid(grid& other) = default;
    grid& operator=(const grid& other) = default;

    // linear index = z*num_x*num_y + y*num_x + x (grid.cpp:161-174)
    Eigen::Vector3i ind2sub(int linear_index) {
        Eigen::Vector3i subs;
        subs[2] = linear_index / (dims[0] * dims[1]);
        linear_index = linear_index % (dims[0] * dims[1]);
        subs[1] = linear_index / dims[0];
        subs[0] = linear_index % dims[0];
        return subs;
    }
    int sub2ind(const Eigen::Vector3i& subs) { return subs[2] * dims[0] * dims[1] + subs[1] * dims[0] + subs[0]; }
    float& operator()(int index) {
        if (index < 0 || index >= dims[0] * dims[1] * dims[2]) {
            std::cerr << "linear index out of range [0:" << dims[0] * dims[1] * dims[2] << "): " << index << std::endl;
            std::exit(1);
        }
        Eigen::Vector3i s = ind2sub(index);
        return voxels[((size_t)s[0] * dims[1] + s[1]) * dims[2] + s[2]];
    }
    Eigen::Vector3f getCloudPoint(const Eigen::Vector3f& pnt) {      // grid.cpp:204-210
        Eigen::Vector3f p;
        for (int a = 0; a < 3; a++) p[a] = ((pnt[a] - (float)pad) * scale[a]) + shift[a];
        return p;
    }
    // flat access for the C-ABI (the storage already is the C-ABI layout: no flatten/unflatten copies)
    float* data() { return voxels.data(); }
    size_t size() const { return voxels.size(); }
};
typedef std::shared_ptr<grid> gridPtr;

inline gridPtr copyGrid(gridPtr in) { return gridPtr(new grid(*in)); }                       // grid.cpp:337-340
inline gridPtr mrb_like(gridPtr g) { return gridPtr(new grid(g->dims, g->scale, g->shift, g->pad)); }
inline gridPtr addGrids(gridPtr a, gridPtr b) {                                             // grid.cpp:343-347
    gridPtr g = mrb_like(a);
    for (size_t i = 0; i < g->size(); i++) g->data()[i] = a->data()[i] + b->data()[i];
    return g;
}
inline gridPtr getBinaryVolume(gridPtr c) {                                                 // grid.cpp:322-334
    gridPtr g = mrb_like(c);
    for (size_t i = 0; i < g->size(); i++) g->data()[i] = c->data()[i] >= 0 ? 1.0f : 0.0f;
    return g;
}
inline std::vector<int> findIndexes(gridPtr band) {                                         // grid.cpp:350-371
    int n = 0;
    mrbcompat::check(mrb_find_indexes(mrbcompat::ctx(), band->data(), &band->dims[0], nullptr, 0, &n));
    std::vector<int> idx((size_t)n, 0);
    if (n) mrbcompat::check(mrb_find_indexes(mrbcompat::ctx(), band->data(), &band->dims[0], idx.data(), n, &n));
    return idx;
}
inline gridPtr getIndexMap(gridPtr band, const std::vector<int>& indexes) {                 // grid.cpp:374-389
    gridPtr m = mrb_like(band);
    mrbcompat::check(mrb_index_map(mrbcompat::ctx(), &band->dims[0], indexes.data(), (int)indexes.size(), m->data()));
    return m;
}

// ---- dfields (dfields.h:13-22,69) ----------------------------------------------------------------------
#define MRB_GRID_OP(NAME, CALL)                                                     \
    inline gridPtr NAME(gridPtr g) {                                                \
        gridPtr out = mrb_like(g);                                                  \
        mrbcompat::check(CALL(mrbcompat::ctx(), g->data(), &g->dims[0], out->data())); \
        return out;                                                                 \
    }
MRB_GRID_OP(fastPerim, mrb_fast_perim)
MRB_GRID_OP(getsqdist, mrb_sqdist)
MRB_GRID_OP(getsqrt, mrb_sqrt)
MRB_GRID_OP(dfield_gpu, mrb_dfield)
// feature pre-pass (dfields.h:49-67)
inline std::vector<int> getSurface(gridPtr volume) {
    int n = 0;
    mrbcompat::check(mrb_get_surface(mrbcompat::ctx(), volume->data(), &volume->dims[0], nullptr, 0, &n));
    std::vector<int> idx((size_t)n, 0);
    if (n) mrbcompat::check(mrb_get_surface(mrbcompat::ctx(), volume->data(), &volume->dims[0], idx.data(), n, &n));
    return idx;
}
inline std::vector<Eigen::Vector3f> getSurfaceNormals(gridPtr volume, const std::vector<int>& surface) {
    std::vector<Eigen::Vector3f> normals(surface.size());
    static_assert(sizeof(Eigen::Vector3f) == 3 * sizeof(float), "Vector3f stand-in must be 3 packed floats");
    mrbcompat::check(mrb_surface_normals(mrbcompat::ctx(), volume->data(), &volume->dims[0], surface.data(), (int)surface.size(),
                                         surface.empty() ? nullptr : &normals[0][0]));
    return normals;
}
inline gridPtr getFeatureMap(gridPtr volume, gridPtr surfaceMap, const std::vector<Eigen::Vector3f>& normals, float threshold) {
    gridPtr fm = mrb_like(volume);
    mrbcompat::check(mrb_feature_map(mrbcompat::ctx(), volume->data(), surfaceMap->data(), normals.empty() ? nullptr : &normals[0][0],
                                     (int)normals.size(), &volume->dims[0], threshold, fm->data()));
    return fm;
}
// ---- narrowBand (narrowBand.h:11-24) ---------------------------------------------------------------------
MRB_GRID_OP(erode_grid, mrb_erode)
MRB_GRID_OP(dilate_grid, mrb_dilate)
#undef MRB_GRID_OP

struct bands {
    gridPtr band;
    gridPtr tight_band;
};
typedef std::shared_ptr<bands> bandsPtr;
inline bandsPtr createBands(gridPtr margin, float band_size) {
    bandsPtr b(new bands());
    b->band = mrb_like(margin);
    b->tight_band = mrb_like(margin);
    mrbcompat::check(mrb_create_bands(mrbcompat::ctx(), margin->data(), &margin->dims[0], band_size, b->band->data(), b->tight_band->data()));
    return b;
}

// ---- primeqp / quadprog (primeqp.h:17-43, quadprog.h:13-26) ----------------------------------------------
// R = H - diag(H) in compressed-column form; stands in for Eigen::SparseMatrix<float>.
struct SparseMatrixCSC {
    int n = 0;
    std::vector<int> jc, ir;
    std::vector<float> pr;
    int rows() const { return n; }
    int cols() const { return n; }
    int nonZeros() const { return (int)ir.size(); }
};
typedef std::shared_ptr<SparseMatrixCSC> SparseMatrixPtr;

struct qp_args {
    SparseMatrixPtr R;
    std::vector<float> invdg, lb, ub, x;
    int iter;
};
typedef std::shared_ptr<qp_args> qp_argsPtr;

inline qp_argsPtr primeQP(gridPtr volume, gridPtr margin, bandsPtr bnds) {
    mrb_ctx* c = mrbcompat::ctx();
    int n = 0, nnz = 0;
    int* dims = &volume->dims[0];
    mrbcompat::check(mrb_find_indexes(c, bnds->band->data(), dims, nullptr, 0, &n));
    qp_argsPtr out(new qp_args());
    out->R.reset(new SparseMatrixCSC());
    out->R->n = n;
    out->invdg.resize(n); out->lb.resize(n); out->ub.resize(n); out->x.resize(n);
    out->R->jc.assign((size_t)n + 1, 0);
    mrbcompat::check(mrb_prime_qp(c, volume->data(), margin->data(), bnds->band->data(), bnds->tight_band->data(), dims, n,
                                  out->lb.data(), out->ub.data(), out->x.data(), out->invdg.data(), out->R->jc.data(),
                                  nullptr, nullptr, 0, &nnz));
    out->R->ir.resize(nnz); out->R->pr.resize(nnz);
    if (nnz)
        mrbcompat::check(mrb_prime_qp(c, volume->data(), margin->data(), bnds->band->data(), bnds->tight_band->data(), dims, n,
                                      nullptr, nullptr, nullptr, nullptr, out->R->jc.data(), out->R->ir.data(),
                                      out->R->pr.data(), nnz, &nnz));
    out->iter = MRB_QP_ITERS_DEFAULT;
    std::cout << "quadratic program ready" << std::endl;     // primeqp.cpp:221
    return out;
}
inline std::vector<int> getFeatureIndexes(gridPtr featureMap, gridPtr indexMap) {          // primeqp.cpp:231-244
    std::vector<int> out;
    for (size_t i = 0; i < featureMap->size(); i++)
        if (featureMap->data()[i] == 1.0f) out.push_back((int)indexMap->data()[i]);
    std::sort(out.begin(), out.end());
    return out;
}
inline std::vector<int> getIr(SparseMatrixPtr R) { return R->ir; }
inline std::vector<int> getJc(SparseMatrixPtr R) { return R->jc; }
inline std::vector<float> getPr(SparseMatrixPtr R) { return R->pr; }

inline std::vector<float> runQP(qp_argsPtr args, const std::vector<int>& featureIndexes, const bool USING_FEATURES) {
    int n = args->R->rows();
    std::vector<float> x((size_t)n, 0.0f);
    mrbcompat::check(mrb_run_qp(mrbcompat::ctx(), n, args->R->jc.data(), args->R->ir.data(), args->R->pr.data(), args->invdg.data(),
                                args->lb.data(), args->ub.data(), args->x.data(), args->iter, featureIndexes.data(),
                                (int)featureIndexes.size(), USING_FEATURES ? 1 : 0, x.data()));
    std::cout << "quadratic program finished" << std::endl;  // quadprog.cpp:103
    return x;
}
inline std::vector<float> runQPGPU(qp_argsPtr a, const std::vector<int>& f, const bool u) { return runQP(a, f, u); }

// USING_CUDA is accepted for signature parity; this library has one compute path (the GPU).
inline gridPtr optimize(gridPtr volume, gridPtr featureMap, const bool USING_FEATURES, const bool USING_CUDA) {
    (void)USING_CUDA;
    gridPtr F = mrb_like(volume);
    mrbcompat::check(mrb_optimize(mrbcompat::ctx(), volume->data(), USING_FEATURES && featureMap ? featureMap->data() : nullptr,
                                  &volume->dims[0], USING_FEATURES ? 1 : 0, MRB_BAND_SIZE_DEFAULT, MRB_QP_ITERS_DEFAULT,
                                  F->data(), nullptr));
    std::cout << "margin calculated" << std::endl << "bands created" << std::endl << "indexes stored" << std::endl;
    std::cout << "quadratic program ready" << std::endl << "quadratic program finished" << std::endl;
    return F;
}

// ---- hasher / mcubes (hasher.h:17-106, mcubes.h:354) --------------------------------------------------------
typedef unsigned long hash_id;
class XYZ {
  public:
    float x, y, z;
    XYZ() {}
    XYZ(float x_, float y_, float z_) : x(x_), y(y_), z(z_) {}
    friend bool operator<(const XYZ& l, const XYZ& r) {
        if (l.x != r.x) return l.x < r.x;
        if (l.y != r.y) return l.y < r.y;
        return l.z < r.z;
    }
};
typedef struct { XYZ p[3]; } TRIANGLE;

class pointHashmap {      // vertex ids in first-appearance order, keyed by hashXYZ (hasher.cpp:10-32)
  public:
    std::vector<XYZ> points;
    int numPoints = 0;
    pointHashmap(std::vector<TRIANGLE> triangles, gridPtr g) : dims_(g->dims) {
        size_t n = triangles.size() * 3;
        std::vector<hash_id> k = keys(n ? &triangles[0].p[0].x : nullptr, (int)n);
        for (size_t i = 0; i < n; i++) {
            bool seen = false;
            for (auto& e : table_[k[i] % kBuckets]) if (e.first == k[i]) { seen = true; break; }
            if (!seen) {
                table_[k[i] % kBuckets].push_back(std::make_pair(k[i], numPoints));
                points.push_back(triangles[i / 3].p[i % 3]);
                numPoints++;
            }
        }
    }
    int index(const XYZ& pnt) {
        hash_id id = keys(&pnt.x, 1)[0];
        for (auto& e : table_[id % kBuckets]) if (e.first == id) return e.second;
        return -1;
    }
  private:
    static const size_t kBuckets = 1u << 16;
    Eigen::Vector3i dims_;
    std::vector<std::pair<hash_id, int> > table_[kBuckets];
    std::vector<hash_id> keys(const float* xyz, int n) {
        std::vector<uint64_t> k((size_t)n);
        if (n) mrbcompat::check(mrb_hash_points(mrbcompat::ctx(), xyz, n, &dims_[0], k.data()));
        return std::vector<hash_id>(k.begin(), k.end());
    }
};

// surfaceMap / normals / thresholds are only consumed when USING_FEATURES (mcubes.cpp:176-183, 234-311, 396-588).
inline void mcubes(gridPtr in, gridPtr surfaceMap, const std::vector<Eigen::Vector3f>& normals, float isolevel,
                   float feat_thresh, float corner_thresh, const char* filename, const bool USING_FEATURES) {
    mrb_mesh mesh;
    if (USING_FEATURES)
        mrbcompat::check(mrb_mcubes_features(mrbcompat::ctx(), in->data(), surfaceMap->data(), normals.empty() ? nullptr : &normals[0][0],
                                             (int)normals.size(), &in->dims[0], isolevel, feat_thresh, corner_thresh, 0, &mesh));
    else
        mrbcompat::check(mrb_mcubes(mrbcompat::ctx(), in->data(), &in->dims[0], isolevel, 0, &mesh));
    mrbcompat::check(mrb_write_ply(filename, mesh.verts, mesh.nverts, mesh.faces, mesh.nfaces, &in->scale[0], &in->shift[0], in->pad));
    mrb_mesh_free(&mesh);
}
