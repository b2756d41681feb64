// frontend.h -- PCL-free restatement of the reference front end (point clouds -> occupancy grid).
//
// SURVEY.md 8(f) row f1 ("next"): needed only so that the mesh_reconstruction CLI accepts the reference's
// inputs (binvox completion + pcd partial view).  It follows mesh_reconstruction.cpp:99-120:
//   getResolutionFactor / binvoxToPCL (binvoxToPCL/binvoxToPcl.cpp:15-219), full_confidence /
//   assign_confidence with ConstConf(1) (confidence_pcl/), voxelizeData (voxelize/voxelize.cpp:6-101),
//   grid(cloud, vox) + fillGrid + getBinaryVolume (grid/grid.cpp:84-121, 213-265, 322-334).
// PARITY UNPINNED: pcl::KdTreeFLANN, pcl::VoxelGrid and pcl::io::loadPCDFile are not in /root/reference; their
// documented behaviour is restated (exact nearest neighbour by a hash grid; VoxelGrid leaf layout
// min_b = floor(min/leaf), divisions = floor(max/leaf) - min_b + 1, one centroid per occupied leaf).
// With ConstConf(1) every strength is 1, so only occupancy survives (grid.cpp:328) and the octree strength
// averaging (voxelize.cpp:54-85) is a no-op that is not restated.
#pragma once

#include <algorithm>
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <fstream>
#include <iostream>
#include <sstream>
#include <string>
#include <unordered_map>
#include <vector>

namespace mrbfront {

struct P3 { float x, y, z; };

struct Binvox {
    int version = 0, depth = -1, height = 0, width = 0;
    float tx = 0, ty = 0, tz = 0, scale = 1;
    std::vector<unsigned char> voxels;
};

// read_binvox, binvoxToPcl.cpp:15-103
inline bool read_binvox(Binvox& vox, const std::string& path) {
    std::ifstream in(path.c_str(), std::ios::in | std::ios::binary);
    std::string line;
    in >> line;
    if (line != "#binvox") { std::cout << "Error: first line reads [" << line << "] instead of [#binvox]" << std::endl; return false; }
    in >> vox.version;
    std::cout << "reading binvox version " << vox.version << std::endl;
    bool done = false;
    while (in.good() && !done) {
        in >> line;
        if (line == "data") done = true;
        else if (line == "dim") in >> vox.depth >> vox.height >> vox.width;
        else if (line == "translate") in >> vox.tx >> vox.ty >> vox.tz;
        else if (line == "scale") in >> vox.scale;
        else {
            std::cout << "  unrecognized keyword [" << line << "], skipping" << std::endl;
            char c;
            do { c = (char)in.get(); } while (in.good() && c != '\n');
        }
    }
    if (!done) { std::cout << "  error reading header" << std::endl; return false; }
    if (vox.depth == -1) { std::cout << "  missing dimensions in header" << std::endl; return false; }
    size_t size = (size_t)vox.width * vox.height * vox.depth;
    vox.voxels.assign(size, 0);
    in.get();   // the linefeed after "data"
    size_t index = 0;
    int nr = 0;
    while (index < size && in.good()) {
        int value = in.get(), count = in.get();
        if (!in.good()) break;
        if (index + (size_t)count > size) return false;
        for (int i = 0; i < count; i++) vox.voxels[index + i] = (unsigned char)value;
        if (value) nr += count;
        index += (size_t)count;
    }
    std::cout << "  read " << nr << " voxels" << std::endl;
    return true;
}

// binvoxToPCL, binvoxToPcl.cpp:105-169
inline std::vector<P3> binvox_to_points(const Binvox& vox, int res) {
    std::vector<P3> cloud;
    const int wxh = vox.width * vox.height;
    for (int x = 0; x < vox.depth - 1; x++)
        for (int y = 0; y < vox.width - 1; y++)
            for (int z = 0; z < vox.height - 1; z++) {
                int idx[8], q = 0;
                for (int i = 0; i <= 1; i++)
                    for (int j = 0; j <= 1; j++)
                        for (int k = 0; k <= 1; k++) idx[q++] = (x + i) * wxh + (z + k) * vox.width + (y + j);
                for (int i = 0; i < res; i++)
                    for (int j = 0; j < res; j++)
                        for (int k = 0; k < res; k++) {
                            float p0 = (float)x + ((float)i) / ((float)res), p1 = (float)y + ((float)j) / ((float)res),
                                  p2 = (float)z + ((float)k) / ((float)res);
                            float val = 0.0f;
                            for (int n = 0; n < 8; n++) {
                                int a = n / 4, b = (n % 4) / 2, c = (n % 4) % 2;
                                float dist = std::fabs(p0 - (float)(x + a));
                                dist += std::fabs(p1 - (float)(y + b));
                                dist += std::fabs(p2 - (float)(z + c));
                                dist += 0.00001;
                                float weight = (float)(2 * vox.voxels[idx[n]] - 1);
                                val += (weight / dist);
                            }
                            if (val > 0.0f) {
                                P3 p;
                                p.x = (p0 + 0.5) * vox.scale / ((float)vox.depth) + vox.tx;
                                p.y = (p1 + 0.5) * vox.scale / ((float)vox.width) + vox.ty;
                                p.z = (p2 + 0.5) * vox.scale / ((float)vox.height) + vox.tz;
                                cloud.push_back(p);
                            }
                        }
            }
    return cloud;
}

// Exact nearest neighbour of the every-10th points among the other points (getDensity, binvoxToPcl.cpp:172-201;
// getResolution, voxelize.cpp:6-32): returns (sum of NN distances, number of queries).
inline void nn_stats(const std::vector<P3>& cloud, double& dist_total, int& count) {
    std::vector<P3> ref;
    for (size_t i = 0; i < cloud.size(); i++) if ((i % 10) > 0) ref.push_back(cloud[i]);
    dist_total = 0; count = 0;
    if (ref.empty()) return;
    float mn[3] = {1e30f, 1e30f, 1e30f}, mx[3] = {-1e30f, -1e30f, -1e30f};
    for (auto& p : ref) { mn[0] = std::min(mn[0], p.x); mn[1] = std::min(mn[1], p.y); mn[2] = std::min(mn[2], p.z);
                          mx[0] = std::max(mx[0], p.x); mx[1] = std::max(mx[1], p.y); mx[2] = std::max(mx[2], p.z); }
    double vol = std::max(1e-30, (double)(mx[0] - mn[0] + 1e-6f) * (mx[1] - mn[1] + 1e-6f) * (mx[2] - mn[2] + 1e-6f));
    float cell = (float)std::cbrt(vol / (double)ref.size()) * 2.0f;
    if (!(cell > 0)) cell = 1.0f;
    auto key = [&](int a, int b, int c) { return ((uint64_t)(uint32_t)a << 42) ^ ((uint64_t)(uint32_t)b << 21) ^ (uint64_t)(uint32_t)c; };
    auto cidx = [&](float v, int ax) { return (int)std::floor((v - mn[ax]) / cell); };
    std::unordered_map<uint64_t, std::vector<int> > cells;
    for (size_t i = 0; i < ref.size(); i++) cells[key(cidx(ref[i].x, 0), cidx(ref[i].y, 1), cidx(ref[i].z, 2))].push_back((int)i);
    float fdist_total = 0.0f;
    for (size_t i = 0; i < cloud.size(); i += 10) {
        const P3& q = cloud[i];
        int c0 = cidx(q.x, 0), c1 = cidx(q.y, 1), c2 = cidx(q.z, 2);
        float best = 1e30f;
        for (int r = 1; r < 1 << 20; r++) {               // grow the search cube until it provably holds the NN
            for (int a = c0 - r; a <= c0 + r; a++)
                for (int b = c1 - r; b <= c1 + r; b++)
                    for (int c = c2 - r; c <= c2 + r; c++) {
                        if (std::max(std::abs(a - c0), std::max(std::abs(b - c1), std::abs(c - c2))) != r && r > 1) continue;
                        auto it = cells.find(key(a, b, c));
                        if (it == cells.end()) continue;
                        for (int id : it->second) {
                            float dx = ref[id].x - q.x, dy = ref[id].y - q.y, dz = ref[id].z - q.z;
                            float d2 = dx * dx + dy * dy + dz * dz;
                            if (d2 < best) best = d2;
                        }
                    }
            if (best < 1e30f && std::sqrt(best) <= (float)(r)*cell) break;
            if (r > 4096) break;
        }
        fdist_total += (float)std::sqrt((double)best);
        count++;
    }
    dist_total = fdist_total;
}

// getResolutionFactor, binvoxToPcl.cpp:203-219
inline int resolution_factor(const Binvox& vox, const std::vector<P3>& observed) {
    double dt; int cnt;
    nn_stats(observed, dt, cnt);
    std::cout << "Total Dist: " << dt << " Count: " << cnt << std::endl;
    float other = 3.0 * ((float)cnt / (float)dt) / 8.0;
    std::vector<P3> init = binvox_to_points(vox, 1);
    nn_stats(init, dt, cnt);
    std::cout << "Total Dist: " << dt << " Count: " << cnt << std::endl;
    float density = (float)cnt / (float)dt;
    std::cout << "Observed density: " << other << std::endl << "Completed density: " << density << std::endl;
    int res = density > other ? 1 : (int)std::ceil(other / density);
    std::cout << "Resolution Factor: " << res << std::endl;
    return res;
}

// loadPCDFile<PointXYZ>: ASCII and uncompressed binary, FIELDS containing x y z (float32).
inline bool load_pcd(const std::string& path, std::vector<P3>& out) {
    std::ifstream in(path.c_str(), std::ios::in | std::ios::binary);
    if (!in.good()) return false;
    std::vector<std::string> fields;
    std::vector<int> sizes, counts;
    long points = -1;
    std::string data_kind, line;
    while (std::getline(in, line)) {
        if (!line.empty() && line[line.size() - 1] == '\r') line.erase(line.size() - 1);
        if (line.empty() || line[0] == '#') continue;
        std::istringstream ss(line);
        std::string k;
        ss >> k;
        if (k == "FIELDS") { std::string f; while (ss >> f) fields.push_back(f); }
        else if (k == "SIZE") { int v; while (ss >> v) sizes.push_back(v); }
        else if (k == "COUNT") { int v; while (ss >> v) counts.push_back(v); }
        else if (k == "POINTS") ss >> points;
        else if (k == "DATA") { ss >> data_kind; break; }
    }
    int ix = -1, iy = -1, iz = -1;
    for (size_t i = 0; i < fields.size(); i++) { if (fields[i] == "x") ix = (int)i; if (fields[i] == "y") iy = (int)i; if (fields[i] == "z") iz = (int)i; }
    if (ix < 0 || iy < 0 || iz < 0 || points < 0) return false;
    if (counts.empty()) counts.assign(fields.size(), 1);
    if (sizes.empty()) sizes.assign(fields.size(), 4);
    out.clear();
    if (data_kind == "ascii") {
        while (std::getline(in, line) && (long)out.size() < points) {
            std::istringstream ss(line);
            std::vector<double> v;
            double t;
            while (ss >> t) v.push_back(t);
            size_t col = 0;
            P3 p = {0, 0, 0};
            bool ok = true;
            for (size_t f = 0; f < fields.size(); f++) {
                if (col >= v.size()) { ok = false; break; }
                if ((int)f == ix) p.x = (float)v[col];
                if ((int)f == iy) p.y = (float)v[col];
                if ((int)f == iz) p.z = (float)v[col];
                col += (size_t)counts[f];
            }
            if (ok && std::isfinite(p.x) && std::isfinite(p.y) && std::isfinite(p.z)) out.push_back(p);
        }
        return true;
    }
    if (data_kind == "binary") {
        size_t stride = 0;
        std::vector<size_t> off(fields.size());
        for (size_t f = 0; f < fields.size(); f++) { off[f] = stride; stride += (size_t)sizes[f] * counts[f]; }
        std::vector<char> rec(stride);
        for (long i = 0; i < points; i++) {
            in.read(rec.data(), (std::streamsize)stride);
            if (!in.good()) break;
            P3 p;
            memcpy(&p.x, &rec[off[ix]], 4); memcpy(&p.y, &rec[off[iy]], 4); memcpy(&p.z, &rec[off[iz]], 4);
            if (std::isfinite(p.x) && std::isfinite(p.y) && std::isfinite(p.z)) out.push_back(p);
        }
        return true;
    }
    return false;   // binary_compressed is not supported
}

struct Volume {
    int dims[3];
    float scale[3], shift[3];
    int pad;
    std::vector<float> voxels;      // [x][y][z] 0/1
};

// voxelizeData + createGrid + getBinaryVolume (voxelize.cpp:6-51,87-101; grid.cpp:84-121,213-265,314-334)
// res_factor <= 0: no fillGrid (the m_cubes executable builds its grid without it, m_cubes.cpp:71)
inline Volume voxelize_leaf(const std::vector<P3>& cloud, float leaf, int res_factor);
inline Volume voxelize(const std::vector<P3>& cloud, int res_factor) {
    double dt; int cnt;
    nn_stats(cloud, dt, cnt);
    float leaf = 2.0 * (float)dt / (float)cnt;
    std::cout << "Leaf size is " << leaf << std::endl;
    return voxelize_leaf(cloud, leaf, res_factor);
}
inline Volume voxelize_leaf(const std::vector<P3>& cloud, float leaf, int res_factor) {
    std::cout << "PointCloud before filtering: " << cloud.size() << " data points" << std::endl;
    float mn[3] = {1e30f, 1e30f, 1e30f}, mx[3] = {-1e30f, -1e30f, -1e30f};
    for (auto& p : cloud) { mn[0] = std::min(mn[0], p.x); mn[1] = std::min(mn[1], p.y); mn[2] = std::min(mn[2], p.z);
                            mx[0] = std::max(mx[0], p.x); mx[1] = std::max(mx[1], p.y); mx[2] = std::max(mx[2], p.z); }
    const float inv = 1.0f / leaf;
    int minb[3], div[3];
    for (int a = 0; a < 3; a++) { minb[a] = (int)std::floor(mn[a] * inv); div[a] = (int)std::floor(mx[a] * inv) - minb[a] + 1; }
    // one centroid per occupied leaf
    std::vector<double> sx((size_t)div[0] * div[1] * div[2], 0.0), sy(sx.size(), 0.0), sz(sx.size(), 0.0);
    std::vector<int> num(sx.size(), 0);
    for (auto& p : cloud) {
        int i = (int)std::floor(p.x * inv) - minb[0], j = (int)std::floor(p.y * inv) - minb[1], k = (int)std::floor(p.z * inv) - minb[2];
        size_t id = ((size_t)i * div[1] + j) * div[2] + k;
        sx[id] += p.x; sy[id] += p.y; sz[id] += p.z; num[id]++;
    }
    Volume g;
    g.pad = 6;
    for (int a = 0; a < 3; a++) g.dims[a] = div[a] + 2 * g.pad;
    std::vector<float> conf((size_t)g.dims[0] * g.dims[1] * g.dims[2], -1.0f);
    float lo[3] = {10000.0f, 10000.0f, 10000.0f}, hi[3] = {-10000.0f, -10000.0f, -10000.0f};
    size_t occupied = 0;
    for (int i = 0; i < div[0]; i++)
        for (int j = 0; j < div[1]; j++)
            for (int k = 0; k < div[2]; k++) {
                size_t id = ((size_t)i * div[1] + j) * div[2] + k;
                if (!num[id]) continue;
                float c[3] = {(float)(sx[id] / num[id]), (float)(sy[id] / num[id]), (float)(sz[id] / num[id])};
                for (int a = 0; a < 3; a++) { lo[a] = std::min(lo[a], c[a]); hi[a] = std::max(hi[a], c[a]); }
                conf[((size_t)(i + g.pad) * g.dims[1] + (j + g.pad)) * g.dims[2] + (k + g.pad)] = 1.0f;
                occupied++;
            }
    std::cout << "PointCloud after filtering: " << occupied << " data points" << std::endl;
    for (int a = 0; a < 3; a++) { g.shift[a] = lo[a]; g.scale[a] = (hi[a] - lo[a]) / ((float)div[a]); }
    // fillGrid (grid.cpp:213-265): close the first gap of every z column if it is narrower than res_factor+1
    const int max_width = res_factor + 1;
    for (int i = 0; i < (res_factor > 0 ? g.dims[0] : 0); i++)
        for (int j = 0; j < g.dims[1]; j++) {
            float* col = &conf[((size_t)i * g.dims[1] + j) * g.dims[2]];
            int num_gap = 0;
            bool first_hit = false, first_gap = false, end_hit = false;
            for (int k = 0; k < g.dims[2] && !end_hit; k++) {
                if (!first_hit) { if (col[k] != -1.0f) first_hit = true; }
                else if (!first_gap) { if (col[k] == -1.0f) { first_gap = true; num_gap++; } }
                else if (col[k] == -1.0f) num_gap++;
                else {
                    end_hit = true;
                    if (num_gap < max_width)
                        for (int n = num_gap; n > 0; n--) {
                            int dist = num_gap - n + 1;
                            float e = -dist * dist / (4 * res_factor);
                            col[k - n] = std::pow(2.71828f, e);
                        }
                }
            }
        }
    g.voxels.resize(conf.size());
    for (size_t q = 0; q < conf.size(); q++) g.voxels[q] = conf[q] >= 0 ? 1.0f : 0.0f;     // getBinaryVolume
    return g;
}

}  // namespace mrbfront
