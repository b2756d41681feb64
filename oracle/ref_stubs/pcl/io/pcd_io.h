// TEST INFRASTRUCTURE: stand-in (binvoxToPcl.h / full_confidence.h include it; the shim hands clouds over in memory, nothing is loaded).
#ifndef REF_STANDIN_PCL_PCD_IO
#define REF_STANDIN_PCL_PCD_IO
#include <string>
#include "pcl/point_types.h"
namespace pcl { namespace io {
template <typename P> int loadPCDFile(const std::string&, PointCloud<P>&) { return -1; }
template <typename P> int savePCDFileASCII(const std::string&, const PointCloud<P>&) { return -1; }
} }
#endif
