// TEST INFRASTRUCTURE: stand-in for the PCL types dragged in by grid.h (PCL is not installed and is not on the hot
// path: SURVEY.md 8c).  Only what grid.cpp / dfields.cpp need in order to compile; the cloud constructor and the
// viewers are never called by the tests.
#ifndef REF_STANDIN_PCL_POINT_TYPES
#define REF_STANDIN_PCL_POINT_TYPES
#include "ref_eigen_standin.h"
#include "boost/shared_ptr.hpp"
namespace pcl {
struct PointXYZ { float x, y, z; };
struct PointXYZRGB { float x, y, z; unsigned char r, g, b; };
struct InterestPoint { float x, y, z, strength; };
struct Normal { float normal_x, normal_y, normal_z; float data_n[4]; float curvature; };
template <typename P> class PointCloud {
public:
    typedef boost::shared_ptr<PointCloud<P> > Ptr;
    std::vector<P> points;
    void push_back(const P& p) { points.push_back(p); }
    size_t size() const { return points.size(); }
};
}  // namespace pcl
#endif
