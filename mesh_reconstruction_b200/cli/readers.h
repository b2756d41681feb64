// readers.h -- host file I/O of the CLIs: the binvox reader (read_binvox, binvoxToPCL/binvoxToPcl.cpp:15-103) and a PCD reader
// standing in for pcl::io::loadPCDFile<PointXYZ> (ASCII and uncompressed binary).  Everything the front end COMPUTES from these
// files runs on the GPU (mrb_front_end / mrb_nn_stats / mrb_voxelize_leaf, csrc/mrb_frontend.cu); its CPU restatement is test
// infrastructure and lives in oracle/frontend_oracle.cpp.
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

}  // namespace mrbfront
