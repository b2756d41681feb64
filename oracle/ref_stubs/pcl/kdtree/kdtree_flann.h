// TEST INFRASTRUCTURE: stand-in (grid.cpp includes it and uses nothing from it).
#include "pcl/point_types.h"
