// test_compat.cpp -- drives the hot path through the reference-named C++ API (compat shim), the way
// mesh_reconstruction.cpp:119-134 does, on a synthetic sphere.  Usage: test_compat <n> <radius> <out.ply>
// Prints band / mesh counts; tests/test_gpu_compat.py compares the PLY with the oracle's.
#include <cmath>
#include "grid.h"
#include "dfields.h"
#include "narrowBand.h"
#include "primeqp.h"
#include "quadprog.h"
#include "mcubes.h"

using namespace std;

int main(int argc, char** argv) {
    int n = argc > 1 ? atoi(argv[1]) : 40;
    float radius = argc > 2 ? (float)atof(argv[2]) : 11.0f;
    const char* out = argc > 3 ? argv[3] : "compat.ply";
    Eigen::Vector3i dims(n, n, n);
    Eigen::Vector3f scale(0.01f, 0.01f, 0.01f), shift(-0.2f, 0.1f, 0.3f);
    gridPtr conf(new grid(dims, scale, shift, 6));
    float c = (n - 1) / 2.0f;
    for (int i = 0; i < n; i++)
        for (int j = 0; j < n; j++)
            for (int k = 0; k < n; k++) {
                float d2 = (i - c) * (i - c) + (j - c) * (j - c) + (k - c) * (k - c);
                (*conf)[i][j][k] = d2 <= radius * radius ? 1.0f : -1.0f;
            }
    gridPtr volume = getBinaryVolume(conf);

    // module-level calls, as optimize() does in quadprog.cu:236-280
    gridPtr margin = getsqrt(getsqdist(fastPerim(volume)));
    gridPtr margin2 = dfield_gpu(volume);
    for (size_t i = 0; i < margin->size(); i++)
        if (margin->data()[i] != margin2->data()[i]) { cerr << "dfield_gpu != getsqrt(getsqdist(fastPerim))" << endl; return 2; }
    bandsPtr bnds = createBands(margin, 4);
    vector<int> indexes = findIndexes(bnds->band);
    gridPtr indexMap = getIndexMap(bnds->band, indexes);
    qp_argsPtr args = primeQP(volume, margin, bnds);
    vector<int> none;
    vector<float> x = runQP(args, none, false);

    gridPtr F = copyGrid(volume);
    for (size_t i = 0; i < F->size(); i++) F->data()[i] = (float)((2.0 * F->data()[i] - 1.0) * (4 + 1.0));
    for (size_t i = 0; i < indexes.size(); i++) (*F)(indexes[i]) = x[i];

    // fused call must agree bit for bit with the module-level composition
    gridPtr F2 = optimize(volume, gridPtr(), false, true);
    for (size_t i = 0; i < F->size(); i++)
        if (F->data()[i] != F2->data()[i]) { cerr << "optimize() != module-level composition at " << i << endl; return 3; }

    // feature pre-pass exactly as mesh_reconstruction.cpp:125-128 (always computed there)
    vector<int> surface = getSurface(volume);
    gridPtr surfaceMap = getIndexMap(volume, surface);
    vector<Eigen::Vector3f> normals = getSurfaceNormals(volume, surface);
    gridPtr featureMap = getFeatureMap(volume, surfaceMap, normals, 0.75f);
    double nfeat = 0;
    for (size_t i = 0; i < featureMap->size(); i++) nfeat += featureMap->data()[i];
    cout << "surface " << surface.size() << " features " << nfeat << endl;
    mcubes(F, surfaceMap, normals, 0.0, 0.75f, 0.8f, out, false);
    cout << "band " << indexes.size() << " nnz " << args->R->nonZeros() << " idmap0 " << (*indexMap)(indexes[0]) << endl;
    return 0;
}
