// TEST INFRASTRUCTURE: stand-in for pcl::octree::OctreePointCloudSearch (not installed, not in the reference tree).
// voxelSearch returns the input points in the cube of edge `resolution` (aligned to multiples of it) that holds the query, or the
// nearest input point if that cube is empty.  modifyStrengths (voxelize.cpp:70-85) only averages the strengths of what comes
// back and every strength is 1 (ConstConf(1), mesh_reconstruction.cpp:107), so only "at least one point" matters; PCL's
// own octree alignment is NOT pinned.
#ifndef REF_STANDIN_PCL_OCTREE
#define REF_STANDIN_PCL_OCTREE
#include <cmath>
#include <vector>
#include "pcl/point_types.h"
namespace pcl { namespace octree {
template <typename P> class OctreePointCloudSearch {
    float res_;
    typename PointCloud<P>::Ptr cloud_;
public:
    typedef boost::shared_ptr<OctreePointCloudSearch<P> > Ptr;
    explicit OctreePointCloudSearch(double resolution) : res_((float)resolution) {}
    void setInputCloud(const typename PointCloud<P>::Ptr& c) { cloud_ = c; }
    void addPointsFromInputCloud() {}
    bool voxelSearch(const P& q, std::vector<int>& out) const {
        out.clear();
        const float c[3] = {std::floor(q.x / res_), std::floor(q.y / res_), std::floor(q.z / res_)};
        int nearest = -1;
        float best = 0.0f;
        for (size_t t = 0; t < cloud_->points.size(); t++) {
            const P& p = cloud_->points[t];
            if (std::floor(p.x / res_) == c[0] && std::floor(p.y / res_) == c[1] && std::floor(p.z / res_) == c[2]) out.push_back((int)t);
            const float d = (p.x - q.x) * (p.x - q.x) + (p.y - q.y) * (p.y - q.y) + (p.z - q.z) * (p.z - q.z);
            if (nearest < 0 || d < best) { best = d; nearest = (int)t; }
        }
        if (out.empty() && nearest >= 0) out.push_back(nearest);
        return !out.empty();
    }
};
} }
#endif
