// TEST INFRASTRUCTURE: stand-in for pcl::KdTreeFLANN (PCL / FLANN are not installed and not in the reference tree).
// nearestKSearch(k = 1) is an EXACT search here (brute force); the squared distance is formed as FLANN's L2_Simple functor
// does, result += diff * diff over x, y, z in float.  What this does not pin: FLANN's own traversal (also exact for the
// default parameters pcl uses, eps = 0), hence nothing but the tie-breaking index, which the callers do not use.
#ifndef REF_STANDIN_PCL_KDTREE_FLANN
#define REF_STANDIN_PCL_KDTREE_FLANN
#include <vector>
#include "pcl/point_types.h"
namespace pcl {
template <typename P> class KdTreeFLANN {
    typename PointCloud<P>::Ptr cloud_;
public:
    void setInputCloud(const typename PointCloud<P>::Ptr& c) { cloud_ = c; }
    int nearestKSearch(const P& q, int k, std::vector<int>& idx, std::vector<float>& d2) const {
        if (k != 1 || !cloud_ || cloud_->points.empty()) return 0;
        float best = 0.0f;
        int bi = -1;
        const size_t n = cloud_->points.size();
        for (size_t t = 0; t < n; t++) {
            const P& r = cloud_->points[t];
            float s = 0.0f, d;
            d = q.x - r.x; s += d * d;
            d = q.y - r.y; s += d * d;
            d = q.z - r.z; s += d * d;
            if (bi < 0 || s < best) { best = s; bi = (int)t; }
        }
        idx.resize(1); d2.resize(1);
        idx[0] = bi; d2[0] = best;
        return 1;
    }
};
}  // namespace pcl
#endif
