// m_cubes -- standard marching cubes of a voxelised point cloud, the reference's second executable (m_cubes.cpp:18-81):
//   m_cubes <pcd filename> <output filename (.ply)> [leaf_size]
// The cloud is voxelised at twice the given (or estimated: mean nearest-neighbour distance) resolution, the occupancy
// grid is meshed at isolevel 0.5 by the GPU marching cubes of libmrb.so and written as ASCII PLY.  Front end: the
// PCL-free restatement in frontend.h (parity unpinned).
#include <cstdlib>
#include <cstring>
#include <iostream>
#include <string>
#include <vector>

#include "frontend.h"
#include "mesh_reconstruction_compat.h"

using namespace std;

int main(int argc, char** argv) {
    const float FEATURE_THRESHOLD = 0.75f, CORNER_THRESHOLD = 0.8f;      // not used without features (m_cubes.cpp:21-22)
    if (argc != 3 && argc != 4) {
        cerr << "Usage: <pcd filename> <output filename (.ply)> [optional]leaf_size" << endl;
        return 1;
    }
    cout << "Running Standard Marching Cubes" << endl;
    const string PCD_PATH(argv[1]), OUT_PATH(argv[2]);
    vector<mrbfront::P3> cloud;
    if (!mrbfront::load_pcd(PCD_PATH, cloud)) {
        cerr << "Couldn't read pcd file " << endl;
        exit(1);
    }
    float resolution;
    if (argc == 4) {
        resolution = (float)atof(argv[3]);
    } else {                                   // getResolution, voxelize.cpp:6-32
        double dt; int cnt;
        mrbfront::nn_stats(cloud, dt, cnt);
        resolution = (float)dt / (float)cnt;
    }
    cout << "Leaf size is " << resolution << endl;
    mrbfront::Volume vol = mrbfront::voxelize_leaf(cloud, 2 * resolution, 0);
    gridPtr volume(new grid(Eigen::Vector3i(vol.dims[0], vol.dims[1], vol.dims[2]), Eigen::Vector3f(vol.scale[0], vol.scale[1], vol.scale[2]),
                            Eigen::Vector3f(vol.shift[0], vol.shift[1], vol.shift[2]), vol.pad));
    memcpy(volume->data(), vol.voxels.data(), vol.voxels.size() * sizeof(float));
    vector<int> surface = getSurface(volume);
    gridPtr surfaceMap = getIndexMap(volume, surface);
    vector<Eigen::Vector3f> normals = getSurfaceNormals(volume, surface);
    mcubes(volume, surfaceMap, normals, 0.5f, FEATURE_THRESHOLD, CORNER_THRESHOLD, OUT_PATH.c_str(), false);
    return 0;
}
