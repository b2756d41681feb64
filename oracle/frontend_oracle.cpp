// frontend_oracle.cpp -- TEST INFRASTRUCTURE: CPU restatement of the reference front end (row f1 of SURVEY.md 8), the
// checker for mesh_reconstruction_b200/csrc/mrb_frontend.cu.  Only tests/, __graft_entry__.smoke() and bench.py's CPU legs
// may call it; the product never does.
//
// Follows mesh_reconstruction.cpp:99-120.  Pinned against the reference's own compiled code (oracle/_ref/libmrb_ref_front.so,
// tests/test_ref_front.py) for everything the reference spells out itself: the up-sampling vote and point transform
// (binvoxToPCL/binvoxToPcl.cpp:105-169), the density / leaf-size sums (binvoxToPcl.cpp:172-219, voxelize/voxelize.cpp:6-32),
// the grid constructor, fillGrid and getBinaryVolume (grid/grid.cpp:84-121,213-265,322-334).
// PARITY UNPINNED for what happens inside PCL, whose source is not in /root/reference (PCL >= 1.2, CMakeLists.txt:7):
//   pcl::KdTreeFLANN::nearestKSearch  -> exact nearest neighbour, squared distance ((dx^2 + dy^2) + dz^2) in float
//   pcl::VoxelGrid::filter            -> leaf (i,j,k) = floor(p * inv_leaf) - floor(min * inv_leaf), divisions = max_b - min_b + 1,
//                                        centroid = float sum of the leaf's points in cloud order / count
//   octree voxelSearch strength mean  -> 1 for every leaf (ConstConf(1): all strengths are 1)
#include <cmath>
#include <cstdint>
#include <cstring>
#include <vector>

namespace {

struct Pt { float x, y, z; };

// ---- binvoxToPCL ------------------------------------------------------------------------------------------------
void upsample(const unsigned char* vox, const int dwh[3], const float xf[4], int res, std::vector<Pt>& out) {
    const int D = dwh[0], W = dwh[1], H = dwh[2];
    const float fres = (float)res;
    const size_t WH = (size_t)W * H;
    out.clear();
    std::vector<float> off((size_t)res);
    for (int i = 0; i < res; i++) off[(size_t)i] = ((float)i) / fres;
    for (int x = 0; x + 1 < D; x++)
        for (int y = 0; y + 1 < W; y++)
            for (int z = 0; z + 1 < H; z++) {
                float w[8];                                  // corner weights, n = 4a + 2b + c
                for (int n = 0; n < 8; n++) {
                    const int a = n >> 2, b = (n >> 1) & 1, c = n & 1;
                    w[n] = (float)(2 * (int)vox[(size_t)(x + a) * WH + (size_t)(z + c) * W + (size_t)(y + b)] - 1);
                }
                for (int i = 0; i < res; i++) {
                    const float px = (float)x + off[(size_t)i];
                    const float ax[2] = {std::fabs(px - (float)x), std::fabs(px - (float)(x + 1))};
                    for (int j = 0; j < res; j++) {
                        const float py = (float)y + off[(size_t)j];
                        const float ay[2] = {std::fabs(py - (float)y), std::fabs(py - (float)(y + 1))};
                        for (int k = 0; k < res; k++) {
                            const float pz = (float)z + off[(size_t)k];
                            const float az[2] = {std::fabs(pz - (float)z), std::fabs(pz - (float)(z + 1))};
                            float val = 0.0f;
                            for (int n = 0; n < 8; n++) {
                                float dist = ax[n >> 2];
                                dist += ay[(n >> 1) & 1];
                                dist += az[n & 1];
                                dist = (float)((double)dist + 0.00001);
                                val += w[n] / dist;
                            }
                            if (!(val > 0.0f)) continue;
                            Pt p;
                            p.x = (float)(((double)px + 0.5) * (double)xf[0] / (double)((float)D) + (double)xf[1]);
                            p.y = (float)(((double)py + 0.5) * (double)xf[0] / (double)((float)W) + (double)xf[2]);
                            p.z = (float)(((double)pz + 0.5) * (double)xf[0] / (double)((float)H) + (double)xf[3]);
                            out.push_back(p);
                        }
                    }
                }
            }
}

// ---- getDensity / getResolution: the nearest-neighbour statistic ------------------------------------------------------
void nn_stats(const Pt* cloud, int n, float& dist_total, int& count) {
    dist_total = 0.0f; count = 0;
    std::vector<Pt> ref;
    for (int i = 0; i < n; i++) if (i % 10) ref.push_back(cloud[i]);
    if (ref.empty()) return;
    const long nq = (n + 9) / 10;
    std::vector<float> best((size_t)nq);
#pragma omp parallel for schedule(static)
    for (long q = 0; q < nq; q++) {
        const Pt& c = cloud[q * 10];
        float b = 1e30f;
        for (size_t t = 0; t < ref.size(); t++) {
            const float dx = ref[t].x - c.x, dy = ref[t].y - c.y, dz = ref[t].z - c.z;
            const float d2 = dx * dx + dy * dy + dz * dz;
            if (d2 < b) b = d2;
        }
        best[(size_t)q] = b;
    }
    for (long q = 0; q < nq; q++) { dist_total += (float)std::sqrt((double)best[(size_t)q]); count++; }
}

// ---- voxelizeData + grid(cloud, vox) + fillGrid + getBinaryVolume ------------------------------------------------------
struct Vol {
    int dims[3], pad, nleaf;
    float scale[3], shift[3];
    std::vector<float> vox;
};

bool voxelize_leaf(const Pt* cloud, int n, float leaf, int res_factor, Vol& g) {
    if (n <= 0 || !(leaf > 0.0f)) return false;
    float mn[3] = {cloud[0].x, cloud[0].y, cloud[0].z}, mx[3] = {cloud[0].x, cloud[0].y, cloud[0].z};
    for (int q = 1; q < n; q++) {
        const float c[3] = {cloud[q].x, cloud[q].y, cloud[q].z};
        for (int a = 0; a < 3; a++) { if (c[a] < mn[a]) mn[a] = c[a]; if (c[a] > mx[a]) mx[a] = c[a]; }
    }
    const float inv = 1.0f / leaf;
    int minb[3], div[3];
    double cells = 1.0, total = 1.0;
    g.pad = 6;
    for (int a = 0; a < 3; a++) {
        minb[a] = (int)std::floor(mn[a] * inv);
        div[a] = (int)std::floor(mx[a] * inv) - minb[a] + 1;
        g.dims[a] = div[a] + 2 * g.pad;
        cells *= div[a]; total *= g.dims[a];
    }
    if (cells > 4294967295.0 || total > 2147483647.0) return false;
    // per leaf: float sums in cloud order
    std::vector<float> sum((size_t)cells * 3, 0.0f);
    std::vector<int> num((size_t)cells, 0);
    for (int q = 0; q < n; q++) {
        const int i = (int)std::floor(cloud[q].x * inv) - minb[0], j = (int)std::floor(cloud[q].y * inv) - minb[1],
                  k = (int)std::floor(cloud[q].z * inv) - minb[2];
        const size_t id = (size_t)i + (size_t)div[0] * ((size_t)j + (size_t)div[1] * (size_t)k);
        sum[3 * id] += cloud[q].x; sum[3 * id + 1] += cloud[q].y; sum[3 * id + 2] += cloud[q].z;
        num[id]++;
    }
    std::vector<signed char> occ((size_t)total, 0);
    float lo[3] = {10000.0f, 10000.0f, 10000.0f}, hi[3] = {-10000.0f, -10000.0f, -10000.0f};
    g.nleaf = 0;
    for (size_t id = 0; id < num.size(); id++) {
        if (!num[id]) continue;
        const float m = (float)num[id];
        for (int a = 0; a < 3; a++) {
            const float c = sum[3 * id + a] / m;
            if (c < lo[a]) lo[a] = c;
            if (c > hi[a]) hi[a] = c;
        }
        const int i = (int)(id % (size_t)div[0]), j = (int)((id / (size_t)div[0]) % (size_t)div[1]), k = (int)(id / ((size_t)div[0] * div[1]));
        occ[((size_t)(i + g.pad) * g.dims[1] + (size_t)(j + g.pad)) * g.dims[2] + (size_t)(k + g.pad)] = 1;
        g.nleaf++;
    }
    for (int a = 0; a < 3; a++) { g.shift[a] = lo[a]; g.scale[a] = (hi[a] - lo[a]) / ((float)div[a]); }
    // fillGrid: per z column, the gap after the first occupied run is closed if it ends and is narrower than res_factor + 1
    // (what is written there is pow(2.71828f, -d*d / (4*res_factor)) > 0, i.e. occupied for getBinaryVolume)
    if (res_factor > 0) {
        const int nz = g.dims[2];
        for (size_t col = 0; col < (size_t)g.dims[0] * g.dims[1]; col++) {
            signed char* c = &occ[col * nz];
            int state = 0, gap_at = -1;
            for (int k = 0; k < nz; k++) {
                if (state == 0) { if (c[k]) state = 1; }
                else if (state == 1) { if (!c[k]) { state = 2; gap_at = k; } }
                else if (c[k]) {
                    if (k - gap_at < res_factor + 1) for (int t = gap_at; t < k; t++) c[t] = 1;
                    break;
                }
            }
        }
    }
    g.vox.resize(occ.size());
    for (size_t q = 0; q < occ.size(); q++) g.vox[q] = occ[q] ? 1.0f : 0.0f;
    return true;
}

}  // namespace

extern "C" {

// returns the number of points; writes up to cap of them
int fo_binvox_to_points(const unsigned char* vox, const int dwh[3], const float xform[4], int res, float* xyz_out, int cap) {
    std::vector<Pt> pts;
    upsample(vox, dwh, xform, res, pts);
    const int n = (int)pts.size();
    if (xyz_out) memcpy(xyz_out, pts.data(), sizeof(Pt) * (size_t)(n < cap ? n : cap));
    return n;
}

void fo_nn_stats(const float* xyz, int n, float* dist_total, int* count) { nn_stats((const Pt*)xyz, n, *dist_total, *count); }

// meta: dims[3], pad, nleaf; fmeta: scale[3], shift[3].  vol_out may be NULL (sizes only).  0 on success.
int fo_voxelize_leaf(const float* xyz, int n, float leaf, int res_factor, int meta[5], float fmeta[6], float* vol_out, long long cap) {
    Vol g;
    if (!voxelize_leaf((const Pt*)xyz, n, leaf, res_factor, g)) return -1;
    for (int a = 0; a < 3; a++) { meta[a] = g.dims[a]; fmeta[a] = g.scale[a]; fmeta[3 + a] = g.shift[a]; }
    meta[3] = g.pad; meta[4] = g.nleaf;
    if (vol_out) {
        if ((long long)g.vox.size() > cap) return -2;
        memcpy(vol_out, g.vox.data(), g.vox.size() * sizeof(float));
    }
    return 0;
}

// mesh_reconstruction.cpp:99-120.  info: res_factor, n_predicted, n_cloud; finfo: observed / completed dist totals, densities, leaf.
int fo_front_end(const unsigned char* vox, const int dwh[3], const float xform[4], const float* observed, int n_obs, int res_in,
                 int meta[5], float fmeta[6], int info[5], float finfo[5], float* vol_out, long long cap) {
    int res = res_in;
    if (res < 1) {
        float dt = 0; int cnt = 0;
        nn_stats((const Pt*)observed, n_obs, dt, cnt);
        finfo[0] = dt; info[3] = cnt;
        const float other = 3.0 * ((float)cnt / dt) / 8.0;
        std::vector<Pt> init;
        upsample(vox, dwh, xform, 1, init);
        nn_stats(init.data(), (int)init.size(), dt, cnt);
        finfo[1] = dt; info[4] = cnt;
        const float density = (float)cnt / dt;
        finfo[2] = other; finfo[3] = density;
        res = density > other ? 1 : (int)std::ceil(other / density);
        if (res < 1) return -3;
    }
    std::vector<Pt> pred;
    upsample(vox, dwh, xform, res, pred);
    std::vector<Pt> all((const Pt*)observed, (const Pt*)observed + n_obs);
    all.insert(all.end(), pred.begin(), pred.end());
    info[0] = res; info[1] = (int)pred.size(); info[2] = (int)all.size();
    float dt = 0; int cnt = 0;
    nn_stats(all.data(), (int)all.size(), dt, cnt);
    const float leaf = 2.0 * dt / (float)cnt;
    finfo[4] = leaf;
    return fo_voxelize_leaf((const float*)all.data(), (int)all.size(), leaf, res, meta, fmeta, vol_out, cap);
}

}  // extern "C"
