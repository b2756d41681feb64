// m_cubes -- standard marching cubes of a voxelised point cloud, the reference's second executable (m_cubes.cpp:18-81):
//   m_cubes <pcd filename> <output filename (.ply)> [leaf_size]
// The cloud is voxelised at twice the given (or estimated: mean nearest-neighbour distance) resolution, the occupancy
// grid is meshed at isolevel 0.5 by the GPU marching cubes of libmrb.so and written as ASCII PLY.  The leaf-size estimate
// and the voxel grid run on the GPU as well (mrb_nn_stats, mrb_voxelize_leaf).
#include <cstdlib>
#include <cstring>
#include <iostream>
#include <string>
#include <vector>

#include "readers.h"
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
        float dt; int cnt;
        mrbcompat::check(mrb_nn_stats(mrbcompat::ctx(), &cloud[0].x, (int)cloud.size(), &dt, &cnt));
        resolution = dt / (float)cnt;
    }
    cout << "Leaf size is " << resolution << endl;
    if (cloud.empty()) { cerr << "empty point cloud" << endl; exit(1); }
    mrb_volume vol;
    mrbcompat::check(mrb_voxelize_leaf(mrbcompat::ctx(), &cloud[0].x, (int)cloud.size(), 2 * resolution, 0, 0, &vol));
    gridPtr volume(new grid(Eigen::Vector3i(vol.dims[0], vol.dims[1], vol.dims[2]), Eigen::Vector3f(vol.scale[0], vol.scale[1], vol.scale[2]),
                            Eigen::Vector3f(vol.shift[0], vol.shift[1], vol.shift[2]), vol.pad));
    memcpy(volume->data(), vol.voxels, (size_t)vol.dims[0] * vol.dims[1] * vol.dims[2] * sizeof(float));
    mrb_volume_free(&vol);
    vector<int> surface = getSurface(volume);
    gridPtr surfaceMap = getIndexMap(volume, surface);
    vector<Eigen::Vector3f> normals = getSurfaceNormals(volume, surface);
    mcubes(volume, surfaceMap, normals, 0.5f, FEATURE_THRESHOLD, CORNER_THRESHOLD, OUT_PATH.c_str(), false);
    return 0;
}
