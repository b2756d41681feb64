// TEST INFRASTRUCTURE: stand-in; the viewers (grid::visualize, visualizeNormals) compile to no-ops.  Graphics: unavailable on this box.
#ifndef REF_STANDIN_PCL_VISUALIZER
#define REF_STANDIN_PCL_VISUALIZER
#include "pcl/point_types.h"
#include "boost/thread/thread.hpp"
namespace pcl { namespace visualization {
enum { PCL_VISUALIZER_POINT_SIZE = 0 };
template <typename P> class PointCloudColorHandlerRGBField {
public:
    explicit PointCloudColorHandlerRGBField(const typename PointCloud<P>::Ptr&) {}
};
class PCLVisualizer {
public:
    explicit PCLVisualizer(const char*) {}
    void setBackgroundColor(double, double, double) {}
    template <typename P> void addPointCloud(const typename PointCloud<P>::Ptr&, const PointCloudColorHandlerRGBField<P>&, const char*) {}
    template <typename P, typename N> void addPointCloudNormals(const typename PointCloud<P>::Ptr&, const typename PointCloud<N>::Ptr&, int, double, const char*) {}
    void setPointCloudRenderingProperties(int, double, const char*) {}
    void addCoordinateSystem(double) {}
    void initCameraParameters() {}
    bool wasStopped() const { return true; }
    void spinOnce(int) {}
};
} }  // namespace pcl::visualization
#endif
