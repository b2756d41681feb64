// TEST INFRASTRUCTURE: stand-in; grid::grid(cloud, vox) (grid.cpp:84-121, front end) compiles against it, nothing calls it.
#ifndef REF_STANDIN_PCL_VOXEL_GRID
#define REF_STANDIN_PCL_VOXEL_GRID
#include "pcl/point_types.h"
namespace pcl {
template <typename P> class VoxelGrid {
public:
    Eigen::Vector3i getMinBoxCoordinates() const { return Eigen::Vector3i(); }
    Eigen::Vector3i getNrDivisions() const { return Eigen::Vector3i(); }
    int getCentroidIndexAt(const Eigen::Vector3i&) const { return -1; }
};
}  // namespace pcl
#endif
