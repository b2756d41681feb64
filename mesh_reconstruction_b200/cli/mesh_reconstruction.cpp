// mesh_reconstruction -- the reference's CLI (mesh_reconstruction.cpp:23-137) on top of libmrb.so.
//
//   mesh_reconstruction <binvox file> <pcd file> <output file> [--feature-detection] [--cuda]
//                       [--feature-threshold f] [--corner-threshold f] [--help]
// Same positional arguments (read raw from argv[1..3], :94-96), same option names, same banner lines
// (:59-82) and the same exit codes (1 on success and on usage errors, 0 for --help; :50,:56,:136).
// --cuda is accepted and selects the only compute path there is: the hot path always runs on the GPU
// (no CPU fallback inside the library).  Extra, not in the reference: --dump-volume <file> writes the
// occupancy grid the front end produced (dims as 3 int32, then float32 voxels) and skips the stages after it.
// The front end itself (mesh_reconstruction.cpp:99-120) also runs on the GPU: mrb_front_end (csrc/mrb_frontend.cu).
#include <cstdlib>
#include <cstring>
#include <iostream>
#include <string>
#include <vector>

#include "../compat/mesh_reconstruction_compat.h"
#include "readers.h"

using namespace std;

static void usage() {
    cout << "Usage: mesh_reconstruction <binvox file> <pcd file> <output file> <options>" << endl;
    cout << "Allowed options:" << endl
         << "  --help                   produce help message" << endl
         << "  --feature-detection      Toggle feature handling" << endl
         << "  --cuda                   Toggle CUDA option" << endl
         << "  --feature-threshold arg  Increasing raises feature sensitivity. Default: 0.75" << endl
         << "  --corner-threshold arg   Decreasing raises feature sensitivity. Default: 0.8" << endl;
}

int main(int argc, char** argv) {
    float FEATURE_THRESHOLD = 0.75f, CORNER_THRESHOLD = 0.8f;
    bool USING_FEATURES = false, USING_CUDA = false, help = false;
    string dump_volume;
    for (int i = 1; i < argc; i++) {
        string a = argv[i];
        if (a == "--help") help = true;
        else if (a == "--feature-detection") USING_FEATURES = true;
        else if (a == "--cuda") USING_CUDA = true;
        else if (a == "--feature-threshold" && i + 1 < argc) FEATURE_THRESHOLD = (float)atof(argv[++i]);
        else if (a.compare(0, 20, "--feature-threshold=") == 0) FEATURE_THRESHOLD = (float)atof(a.c_str() + 20);
        else if (a == "--corner-threshold" && i + 1 < argc) CORNER_THRESHOLD = (float)atof(argv[++i]);
        else if (a.compare(0, 19, "--corner-threshold=") == 0) CORNER_THRESHOLD = (float)atof(a.c_str() + 19);
        else if (a == "--dump-volume" && i + 1 < argc) dump_volume = argv[++i];
        else if (a.compare(0, 2, "--") == 0) { cerr << "error: unrecognised option '" << a << "'" << endl; return 1; }
    }
    if (argc < 4) { usage(); return 1; }
    if (help) { usage(); return 0; }
    cout << (USING_FEATURES ? "Using feature detection" : "Not using feature detection") << endl;
    cout << (USING_CUDA ? "Using CUDA" : "Not using CUDA") << endl;
    cout << "Feature threshold is " << FEATURE_THRESHOLD << endl;
    cout << "Corner threshold is " << CORNER_THRESHOLD << endl;

    string binvox_path(argv[1]), pcd_path(argv[2]), output_path(argv[3]);
    vector<mrbfront::P3> observe;
    if (!mrbfront::load_pcd(pcd_path, observe)) { cerr << "Couldn't read pcd file " << endl; exit(1); }
    mrbfront::Binvox vox;
    if (!mrbfront::read_binvox(vox, binvox_path)) { cout << "Error reading [" << binvox_path << "]" << endl << endl; exit(1); }
    /* getResolutionFactor, binvoxToPCL, full_confidence + assign_confidence(ConstConf(1)), voxelizeData, createGrid and
       getBinaryVolume (mesh_reconstruction.cpp:104-120) in one GPU call; the reference's progress lines from what it reports */
    const int dwh[3] = {vox.depth, vox.width, vox.height};
    const float xform[4] = {vox.scale, vox.tx, vox.ty, vox.tz};
    mrb_volume vol;
    mrb_front_info fi;
    mrbcompat::check(mrb_front_end(mrbcompat::ctx(), vox.voxels.data(), dwh, xform, observe.empty() ? nullptr : &observe[0].x, (int)observe.size(),
                                   0, 0, &vol, &fi));
    cout << "Total Dist: " << fi.observed_dist_total << " Count: " << fi.observed_count << endl;
    cout << "Total Dist: " << fi.completed_dist_total << " Count: " << fi.completed_count << endl;
    cout << "Observed density: " << fi.observed_density << endl << "Completed density: " << fi.completed_density << endl;
    cout << "Resolution Factor: " << fi.res_factor << endl;
    cout << "Leaf size is " << fi.leaf << endl;
    cout << "PointCloud before filtering: " << fi.n_cloud << " data points" << endl;
    cout << "PointCloud after filtering: " << vol.occupied_leaves << " data points" << endl;
    const size_t nvox = (size_t)vol.dims[0] * vol.dims[1] * vol.dims[2];

    if (!dump_volume.empty()) {
        FILE* f = fopen(dump_volume.c_str(), "wb");
        if (!f) { cerr << "cannot write " << dump_volume << endl; return 1; }
        fwrite(vol.dims, sizeof(int), 3, f);
        fwrite(vol.voxels, sizeof(float), nvox, f);
        fclose(f);
        return 1;
    }

    gridPtr volume(new grid(Eigen::Vector3i(vol.dims[0], vol.dims[1], vol.dims[2]), Eigen::Vector3f(vol.scale[0], vol.scale[1], vol.scale[2]),
                            Eigen::Vector3f(vol.shift[0], vol.shift[1], vol.shift[2]), vol.pad));
    memcpy(volume->data(), vol.voxels, nvox * sizeof(float));
    mrb_volume_free(&vol);

    /* Perform feature detection (mesh_reconstruction.cpp:123-128) */
    vector<int> surface = getSurface(volume);
    gridPtr surfaceMap = getIndexMap(volume, surface);
    vector<Eigen::Vector3f> normals = getSurfaceNormals(volume, surface);
    gridPtr featureMap = getFeatureMap(volume, surfaceMap, normals, FEATURE_THRESHOLD);
    /* Perform smoothing */
    gridPtr F = optimize(volume, featureMap, USING_FEATURES, USING_CUDA);
    /* Extract mesh and write to file */
    mcubes(F, surfaceMap, normals, 0.0, FEATURE_THRESHOLD, CORNER_THRESHOLD, output_path.c_str(), USING_FEATURES);
    return 1;
}
