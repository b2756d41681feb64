// TEST INFRASTRUCTURE: stand-in for pcl::VoxelGrid (PCL is not installed and not in the reference tree): the filter as PCL
// documents and implements it (voxel_grid.hpp: applyFilter with save_leaf_layout and downsample_all_data), restated:
//   min_b = floor(min_p * inverse_leaf_size), max_b likewise, div_b = max_b - min_b + 1, divb_mul = (1, div_b.x, div_b.x * div_b.y)
//   per point idx = (floor(p * inverse_leaf_size) - min_b) . divb_mul; points sorted by idx; per run one centroid = sum of all
//   fields in float / count, emitted in idx order; leaf_layout[idx] = position of the centroid in the output
//   getCentroidIndexAt(ijk) = leaf_layout[(ijk - min_b) . divb_mul], -1 when that is out of range
// NOT pinned (inside PCL): the order in which the points of one leaf are summed (std::sort there; cloud order here) and
// PCL-version differences of the centroid type.  grid::grid(cloud, vox) (grid.cpp:84-121), which calls this, IS the reference's code.
#ifndef REF_STANDIN_PCL_VOXEL_GRID
#define REF_STANDIN_PCL_VOXEL_GRID
#include <algorithm>
#include <cmath>
#include <vector>
#include "pcl/point_types.h"
namespace pcl {
template <typename P> class VoxelGrid {
    typename PointCloud<P>::Ptr in_;
    float inv_[3];
    int min_b_[3], max_b_[3], div_b_[3], mul_[3];
    std::vector<int> layout_;
    struct Key { unsigned idx; int pt; };
    static bool key_less(const Key& a, const Key& b) { return a.idx < b.idx; }
public:
    VoxelGrid() { for (int a = 0; a < 3; a++) { inv_[a] = 0; min_b_[a] = max_b_[a] = div_b_[a] = mul_[a] = 0; } }
    void setInputCloud(const typename PointCloud<P>::Ptr& c) { in_ = c; }
    void setLeafSize(float lx, float ly, float lz) { inv_[0] = 1.0f / lx; inv_[1] = 1.0f / ly; inv_[2] = 1.0f / lz; }
    void setSaveLeafLayout(bool) {}
    void setDownsampleAllData(bool) {}
    void filter(PointCloud<P>& out) {
        out.points.clear();
        layout_.clear();
        const std::vector<P>& pts = in_->points;
        if (pts.empty()) return;
        float mn[3] = {pts[0].x, pts[0].y, pts[0].z}, mx[3] = {pts[0].x, pts[0].y, pts[0].z};
        for (size_t q = 1; q < pts.size(); q++) {
            const float c[3] = {pts[q].x, pts[q].y, pts[q].z};
            for (int a = 0; a < 3; a++) { mn[a] = std::min(mn[a], c[a]); mx[a] = std::max(mx[a], c[a]); }
        }
        for (int a = 0; a < 3; a++) {
            min_b_[a] = static_cast<int>(std::floor(mn[a] * inv_[a]));
            max_b_[a] = static_cast<int>(std::floor(mx[a] * inv_[a]));
            div_b_[a] = max_b_[a] - min_b_[a] + 1;
        }
        mul_[0] = 1; mul_[1] = div_b_[0]; mul_[2] = div_b_[0] * div_b_[1];
        std::vector<Key> keys(pts.size());
        for (size_t q = 0; q < pts.size(); q++) {
            const int i = static_cast<int>(std::floor(pts[q].x * inv_[0]) - static_cast<float>(min_b_[0]));
            const int j = static_cast<int>(std::floor(pts[q].y * inv_[1]) - static_cast<float>(min_b_[1]));
            const int k = static_cast<int>(std::floor(pts[q].z * inv_[2]) - static_cast<float>(min_b_[2]));
            keys[q].idx = (unsigned)(i * mul_[0] + j * mul_[1] + k * mul_[2]);
            keys[q].pt = (int)q;
        }
        std::stable_sort(keys.begin(), keys.end(), key_less);
        layout_.assign((size_t)div_b_[0] * div_b_[1] * div_b_[2], -1);
        size_t s = 0;
        while (s < keys.size()) {
            size_t e = s;
            float sum[4] = {0.0f, 0.0f, 0.0f, 0.0f};
            while (e < keys.size() && keys[e].idx == keys[s].idx) {
                const P& p = pts[(size_t)keys[e].pt];
                sum[0] += p.x; sum[1] += p.y; sum[2] += p.z; sum[3] += p.strength;
                e++;
            }
            const float m = static_cast<float>(e - s);
            P c;
            c.x = sum[0] / m; c.y = sum[1] / m; c.z = sum[2] / m; c.strength = sum[3] / m;
            layout_[keys[s].idx] = (int)out.points.size();
            out.points.push_back(c);
            s = e;
        }
    }
    static Eigen::Vector3i v3(const int* a) { Eigen::Vector3i v; v[0] = a[0]; v[1] = a[1]; v[2] = a[2]; return v; }
    Eigen::Vector3i getMinBoxCoordinates() const { return v3(min_b_); }
    Eigen::Vector3i getMaxBoxCoordinates() const { return v3(max_b_); }
    Eigen::Vector3i getNrDivisions() const { return v3(div_b_); }
    int getCentroidIndexAt(const Eigen::Vector3i& ijk) const {
        const int idx = (ijk[0] - min_b_[0]) * mul_[0] + (ijk[1] - min_b_[1]) * mul_[1] + (ijk[2] - min_b_[2]) * mul_[2];
        if (idx < 0 || idx >= static_cast<int>(layout_.size())) return -1;
        return layout_[(size_t)idx];
    }
};
}  // namespace pcl
#endif
