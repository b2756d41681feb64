// ref_front_shim.cpp -- TEST INFRASTRUCTURE.  Flat-array C entry points over the reference's OWN front-end translation units
// (binvoxToPCL/binvoxToPcl.cpp, confidence_pcl/{full_confidence,assign_confidence,ConstConf}.cpp, voxelize/voxelize.cpp,
// grid/grid.cpp), compiled unmodified from /root/reference by build_ref.sh against the PCL stand-ins of ref_stubs/pcl/.
// The calls below are the reference's (mesh_reconstruction.cpp:99-120); only the containers are converted.
#include <cstring>
#include <iostream>
#include <sstream>
#include <vector>

#include "binvoxToPcl.h"
#include "assign_confidence.h"
#include "ConstConf.h"
#include "full_confidence.h"
#include "grid.h"
#include "voxelize.h"

float getDensity(pcl::PointCloud<pcl::PointXYZ>::Ptr cloud);      // binvoxToPcl.cpp:172 (not declared in its header)

namespace {
struct Quiet {                                  // the reference prints progress on cout
    std::streambuf* old;
    std::ostringstream sink;
    Quiet() : old(std::cout.rdbuf(sink.rdbuf())) {}
    ~Quiet() { std::cout.rdbuf(old); }
};
pcl::PointCloud<pcl::PointXYZ>::Ptr to_cloud(const float* xyz, int n) {
    pcl::PointCloud<pcl::PointXYZ>::Ptr c(new pcl::PointCloud<pcl::PointXYZ>());
    for (int i = 0; i < n; i++) { pcl::PointXYZ p; p.x = xyz[3 * i]; p.y = xyz[3 * i + 1]; p.z = xyz[3 * i + 2]; c->push_back(p); }
    return c;
}
int volume_out(gridPtr vol, int meta[4], float fmeta[6], float* out, long long cap) {
    for (int a = 0; a < 3; a++) { meta[a] = vol->dims[a]; fmeta[a] = vol->scale[a]; fmeta[3 + a] = vol->shift[a]; }
    meta[3] = vol->pad;
    if (!out) return 0;
    if ((long long)vol->dims[0] * vol->dims[1] * vol->dims[2] > cap) return -2;
    size_t q = 0;
    for (int i = 0; i < vol->dims[0]; i++)
        for (int j = 0; j < vol->dims[1]; j++)
            for (int k = 0; k < vol->dims[2]; k++) out[q++] = (*vol)[i][j][k];
    return 0;
}
}  // namespace

extern "C" {

// binvoxToPCL(path, res): returns the number of points, writes up to cap.
int reff_binvox_to_points(const char* path, int res, float* xyz_out, int cap) {
    Quiet q;
    pcl::PointCloud<pcl::PointXYZ>::Ptr c = binvoxToPCL(std::string(path), res);
    const int n = (int)c->points.size();
    for (int i = 0; i < n && i < cap && xyz_out; i++) { xyz_out[3 * i] = c->points[i].x; xyz_out[3 * i + 1] = c->points[i].y; xyz_out[3 * i + 2] = c->points[i].z; }
    return n;
}
// read_binvox: header + decoded voxels (vox_out may be NULL); hdr = depth, height, width; fhdr = tx, ty, tz, scale
int reff_read_binvox(const char* path, int hdr[3], float fhdr[4], unsigned char* vox_out, long long cap) {
    Quiet q;
    binvox v;
    if (!read_binvox(&v, std::string(path))) return -1;
    hdr[0] = v.depth; hdr[1] = v.height; hdr[2] = v.width;
    fhdr[0] = v.tx; fhdr[1] = v.ty; fhdr[2] = v.tz; fhdr[3] = v.scale;
    if (vox_out) { if (v.size > cap) return -2; memcpy(vox_out, v.voxels, (size_t)v.size); }
    return v.size;
}
float reff_density(const float* xyz, int n) { Quiet q; return getDensity(to_cloud(xyz, n)); }
int reff_resolution_factor(const char* path, const float* observed, int n) { Quiet q; return getResolutionFactor(std::string(path), to_cloud(observed, n)); }
// getResolution on the cloud with strengths 1
float reff_resolution(const float* xyz, int n) { Quiet q; return getResolution(full_confidence(to_cloud(xyz, n))); }

// voxelizeData + createGrid + getBinaryVolume on a given cloud (all strengths 1) with a given res_factor
int reff_volume_of_cloud(const float* xyz, int n, int res_factor, int meta[4], float fmeta[6], float* leaf_out, float* vol_out, long long cap) {
    Quiet q;
    pcl::PointCloud<pcl::InterestPoint>::Ptr conf = full_confidence(to_cloud(xyz, n));
    voxelized_dataPtr data = voxelizeData(conf);
    if (leaf_out) *leaf_out = data->resolution;
    gridPtr grid_cloud = createGrid(data->filtered_cloud, data->grid_data, res_factor);
    gridPtr volume = getBinaryVolume(grid_cloud);
    return volume_out(volume, meta, fmeta, vol_out, cap);
}

// mesh_reconstruction.cpp:99-120 from the binvox path and the observed points; info: res, n_predicted, n_cloud
int reff_front_end(const char* binvox_path, const float* observed, int n_obs, int meta[4], float fmeta[6], int info[3], float* leaf_out,
                   float* vol_out, long long cap) {
    Quiet q;
    pcl::PointCloud<pcl::PointXYZ>::Ptr observeCloud = to_cloud(observed, n_obs);
    int res = getResolutionFactor(binvox_path, observeCloud);
    pcl::PointCloud<pcl::PointXYZ>::Ptr predictCloud = binvoxToPCL(binvox_path, res);
    Confidencor* confidence_assigner = new ConstConf(1);
    pcl::PointCloud<pcl::InterestPoint>::Ptr confPCL = full_confidence(observeCloud);
    assign_confidence(confPCL, predictCloud, confidence_assigner);
    info[0] = res; info[1] = (int)predictCloud->points.size(); info[2] = (int)confPCL->points.size();
    voxelized_dataPtr data = voxelizeData(confPCL);
    if (leaf_out) *leaf_out = data->resolution;
    gridPtr grid_cloud = createGrid(data->filtered_cloud, data->grid_data, res);
    gridPtr volume = getBinaryVolume(grid_cloud);
    delete confidence_assigner;
    return volume_out(volume, meta, fmeta, vol_out, cap);
}

}  // extern "C"
