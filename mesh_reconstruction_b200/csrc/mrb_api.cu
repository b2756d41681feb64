// mrb_api.cu -- fused device-resident path: optimize (stages 1-3), reconstruct (stages 1-4), PLY writer.
//
// Replaces optimize (optimize/quadprog.cu:236-280 / quadprog.cpp:111-147) and the driver part of
// mcubes (marching_cubes/mcubes.cpp:592-662).  The reference moves the grid host<->device around
// every GPU stage and does perimeter->band->H assembly->F->marching cubes on the CPU even with
// --cuda; here the occupancy grid is uploaded once and only the mesh comes back.
#include "mrb_internal.cuh"

#include <stdlib.h>
#include <string.h>
#include <algorithm>
#include <string>
#include <chrono>
#include <thread>
#include <vector>

namespace mrb {

int export_mesh(mrb_ctx* ctx, float* verts_dev, int* faces_dev, int nverts, int nfaces, int device_out, mrb_mesh* out);

// stages 1-3, implemented in mrb_strips.cu
int optimize_strips(mrb_ctx* ctx, const float* vol, const float* fmap, Dims d, int using_features, int band_size,
                    int iters, float* F_dev, mrb_qp_stats* stats);

static int optimize_dev(mrb_ctx* ctx, const float* vol, const float* fmap, Dims d, int using_features, int band_size,
                        int iters, float* F_dev, mrb_qp_stats* stats) {
    return optimize_strips(ctx, vol, fmap, d, using_features, band_size, iters, F_dev, stats);
}

// Stages 1-3 when marching cubes follows at `isolevel`: the kernel that writes F also packs its sign bits.
static int optimize_for_mc(mrb_ctx* ctx, const float* vol, Dims d, int band_size, int iters, float isolevel, float* F_dev,
                           mrb_qp_stats* stats) {
    ctx->want_sign_bits = true;
    ctx->sign_iso = isolevel;
    ctx->sign_bits = nullptr;            // never a leftover of an earlier call
    const int rc = optimize_strips(ctx, vol, nullptr, d, 0, band_size, iters, F_dev, stats);
    ctx->want_sign_bits = false;
    return rc;
}

}  // namespace mrb

using namespace mrb;

// ---- batch of independent objects ------------------------------------------------------------
// `volumes` holds nobj grids of dims[] one after the other; with x the slowest axis that IS one grid of nobj * nx
// x-planes.  If the first and last band_size + 2 x-planes of every object are empty, nothing an object produces can
// see its neighbours in that stack (perimeter windows, the +-band_size EDT window, erosion / dilation and the radius-2
// QP star all stay inside the object's own planes, and no surface cell spans two objects), so the whole stack goes
// through stages 1-4 ONCE, at large-grid efficiency, instead of about 50 launches and 8 host round trips per 52^3 object.
// Marching cubes computes vertex coordinates and hash keys local to each object (McStack), which makes every
// object's vertices, first-appearance order and faces identical to a run on that object alone.
__global__ void k_quiet_borders(const float* __restrict__ vol, int nx, size_t plane, int nobj, int margin, int* __restrict__ bad) {
    const size_t per_obj = (size_t)2 * margin * plane, total = per_obj * nobj;
    bool any = false;
    for (size_t t = (size_t)blockIdx.x * blockDim.x + threadIdx.x; t < total; t += (size_t)gridDim.x * blockDim.x) {
        const int ob = (int)(t / per_obj);
        const size_t r = t % per_obj;
        int xp = (int)(r / plane);
        if (xp >= margin) xp = nx - 2 * margin + xp;          // the last `margin` planes
        any = any || vol[((size_t)ob * nx + xp) * plane + r % plane] != 0.0f;
    }
    if (any) *bad = 1;
}

// memcpy into freshly malloc'd memory is dominated by first-touch page faults (measured: 90 MB in 38 ms, 2.4 GB/s); several
// threads fault their own slices in parallel.
static void parallel_memcpy(void* dst, const void* src, size_t bytes) {
    const size_t kSlice = (size_t)4 << 20;
    unsigned hw = std::thread::hardware_concurrency();
    size_t nth = std::min<size_t>(std::min<size_t>(hw ? hw : 4, 8), bytes / kSlice);
    if (nth < 2) { memcpy(dst, src, bytes); return; }
    std::vector<std::thread> th;
    for (size_t t = 0; t < nth; t++) {
        const size_t a = bytes * t / nth / 64 * 64, b = t + 1 == nth ? bytes : bytes * (t + 1) / nth / 64 * 64;
        th.emplace_back([=] { memcpy((char*)dst + a, (const char*)src + a, b - a); });
    }
    for (auto& t : th) t.join();
}

// The batch's host mesh: malloc'd arrays that become mrb_mesh.verts / .faces as they are (no second copy), grown by doubling;
// a chunk's result goes device -> pinned staging (full PCIe rate) -> these arrays.
struct HostMesh {
    float* v = nullptr;
    int* f = nullptr;
    size_t nv3 = 0, nf3 = 0, capv = 0, capf = 0;
    double grow_hint = 1.0;          // after the first chunk: (objects of the batch) / (objects of that chunk), so that one allocation holds all
    int append(mrb_ctx* ctx, const float* dverts, int nv, const int* dfaces, int nf) {
        const size_t av = (size_t)nv * 3, af = (size_t)nf * 3;
        if (nv3 + av > capv) {
            size_t c = std::max(nv3 + av, capv * 2);
            if (capv == 0) c = std::max(c, (size_t)((double)av * grow_hint * 1.08) + 1024);
            float* p = (float*)realloc(v, sizeof(float) * (c ? c : 1));
            if (!p) { set_error("mrb_reconstruct_batch: host allocation failed"); return MRB_ENOMEM; }
            v = p; capv = c;
        }
        if (nf3 + af > capf) {
            size_t c = std::max(nf3 + af, capf * 2);
            if (capf == 0) c = std::max(c, (size_t)((double)af * grow_hint * 1.08) + 1024);
            int* p = (int*)realloc(f, sizeof(int) * (c ? c : 1));
            if (!p) { set_error("mrb_reconstruct_batch: host allocation failed"); return MRB_ENOMEM; }
            f = p; capf = c;
        }
        const size_t vb = (sizeof(float) * av + 255) / 256 * 256, fb = sizeof(int) * af;
        char* pin = nullptr;
        MRB_TRY(pinned_reserve(ctx, vb + fb + 256, &pin));
        if (av) MRB_CUDA(cudaMemcpyAsync(pin, dverts, sizeof(float) * av, cudaMemcpyDeviceToHost, ctx->stream));
        if (af) MRB_CUDA(cudaMemcpyAsync(pin + vb, dfaces, fb, cudaMemcpyDeviceToHost, ctx->stream));
        MRB_TRY(sync(ctx));
        if (av) parallel_memcpy(v + nv3, pin, sizeof(float) * av);
        if (af) parallel_memcpy(f + nf3, pin + vb, fb);
        nv3 += av; nf3 += af;
        return MRB_OK;
    }
    void release() { free(v); free(f); v = nullptr; f = nullptr; }
};

static int batch_sequential(mrb_ctx* ctx, const float* volumes, Dims d, int ob0, int ob1, const mrb_options& o,
                            HostMesh& hm, int* vert_off, int* face_off) {
    for (int ob = ob0; ob < ob1; ob++) {
        MRB_TRY(ctx->arena.reset());
        const float* dv = nullptr;
        MRB_TRY(stage_in(ctx, volumes + (size_t)ob * d.count(), d.count(), &dv));
        float* F = nullptr;
        MRB_TRY(arena_alloc(ctx, d.count(), &F));
        int rc = optimize_for_mc(ctx, dv, d, o.band_size, o.iters, o.isolevel, F, nullptr);
        if (rc != MRB_OK) { std::string m = mrb_last_error(); set_error("object %d: %s", ob, m.c_str()); return rc; }
        float* dverts = nullptr;
        int* dfaces = nullptr;
        int nv = 0, nf = 0;
        MRB_TRY(dev_mcubes(ctx, F, d, o.isolevel, &dverts, &dfaces, &nv, &nf));
        MRB_TRY(hm.append(ctx, dverts, nv, dfaces, nf));
        vert_off[ob + 1] = vert_off[ob] + nv;
        face_off[ob + 1] = face_off[ob] + nf;
    }
    return MRB_OK;
}

// One chunk of objects as one stacked grid.  *done = 0 if the chunk does not qualify (the caller runs it object by object).
static int batch_stacked(mrb_ctx* ctx, const float* volumes, Dims d, int ob0, int ob1, const mrb_options& o,
                         HostMesh& hm, int* vert_off, int* face_off, int* done) {
    *done = 0;
    const double t_begin = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now().time_since_epoch()).count();
    const int n = ob1 - ob0, margin = o.band_size + 2;
    if (n < 2 || d.nx < 2 * margin + 1 || !(fabsf(o.isolevel) < (float)(o.band_size + 1))) return MRB_OK;
    // hash keys of different objects must not collide: shift by a power of two above the largest key of one object
    const double kmax = (double)d.nx * 1e9 * (double)(d.ny * d.nz) + (double)d.ny * 1e6 * (double)d.nz + (double)d.nz * 1e3;
    unsigned long long kstride = 1;
    while ((double)kstride <= kmax && kstride < (1ull << 62)) kstride <<= 1;
    if ((double)kstride * (double)n >= 9.0e18) return MRB_OK;
    MRB_TRY(ctx->arena.reset());
    const size_t cnt = d.count() * (size_t)n;
    const float* dv = nullptr;
    MRB_TRY(stage_in(ctx, volumes + (size_t)ob0 * d.count(), cnt, &dv));
    int* bad = ctx->d_mail + 56;
    MRB_CUDA(cudaMemsetAsync(bad, 0, sizeof(int), ctx->stream));
    MRB_LAUNCH(ctx, k_quiet_borders, ctx->sm_count * 8, 256, 0, dv, d.nx, (size_t)d.ny * d.nz, n, margin, bad);
    int is_bad = 0;
    MRB_TRY(read_int(ctx, bad, &is_bad));
    if (is_bad) return MRB_OK;
    Dims ds{d.nx * n, d.ny, d.nz};
    float* F = nullptr;
    MRB_TRY(arena_alloc(ctx, cnt, &F));
    if (optimize_for_mc(ctx, dv, ds, o.band_size, o.iters, o.isolevel, F, nullptr) != MRB_OK) return MRB_OK;   // sequential run names the object
    std::vector<int> foff((size_t)n + 1), voff((size_t)n + 1);
    int seam = 0;
    McStack st{n, d.nx, kstride, foff.data(), voff.data(), &seam};
    float* dverts = nullptr;
    int* dfaces = nullptr;
    int nv = 0, nf = 0;
    MRB_TRY(dev_mcubes(ctx, F, ds, o.isolevel, &dverts, &dfaces, &nv, &nf, &st));
    if (seam) return MRB_OK;
    const bool verbose = getenv("MRB_BATCH_VERBOSE") != nullptr;
    auto now = [] { return std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now().time_since_epoch()).count(); };
    double t_gpu = 0;
    if (verbose) { cudaStreamSynchronize(ctx->stream); t_gpu = now(); }
    MRB_TRY(hm.append(ctx, dverts, nv, dfaces, nf));
    if (verbose) fprintf(stderr, "[mrb] batch chunk of %d objects: stages 1-4 %.2f ms, mesh to host %.2f ms (%d faces)\n", n, t_gpu - t_begin, now() - t_gpu, nf);
    for (int i = 0; i < n; i++) {
        vert_off[ob0 + i + 1] = vert_off[ob0] + voff[(size_t)i + 1];
        face_off[ob0 + i + 1] = face_off[ob0] + foff[(size_t)i + 1];
    }
    *done = 1;
    return MRB_OK;
}

extern "C" {

int mrb_optimize(mrb_ctx* ctx, const float* volume, const float* feature_map, const int dims[3], int using_features,
                 int band_size, int iters, float* F_out, mrb_qp_stats* stats) {
    MRB_REQUIRE(ctx && volume && dims && F_out, "NULL argument");
    MRB_REQUIRE(band_size >= 1 && band_size <= 64 && iters >= 0, "bad band_size / iters");
    MRB_REQUIRE(!using_features || feature_map, "using_features needs a feature map");
    MRB_CUDA(cudaSetDevice(ctx->device));
    MRB_TRY(ctx->arena.reset());
    Dims d = mkdims(dims);
    MRB_REQUIRE(d.count() <= (size_t)0x7fffffff, "grid too large");
    const float *dv = nullptr, *df = nullptr;
    MRB_TRY(stage_in(ctx, volume, d.count(), &dv));
    if (using_features) MRB_TRY(stage_in(ctx, feature_map, d.count(), &df));
    Out<float> oF;
    MRB_TRY(oF.init(ctx, F_out, d.count()));
    MRB_TRY(optimize_dev(ctx, dv, df, d, using_features, band_size, iters, oF.dev, stats));
    MRB_TRY(oF.commit(ctx));
    return sync(ctx);
}

// --feature-detection end to end (mesh_reconstruction.cpp:123-134): getSurface, getIndexMap, getSurfaceNormals,
// getFeatureMap, optimize(USING_FEATURES), mcubes(USING_FEATURES).  Composed from the module-level entry points on
// device buffers that live outside the per-call arena (each entry point resets the arena).
static int reconstruct_with_features(mrb_ctx* ctx, const float* volume, const int dims[3], const mrb_options& o, int device_out,
                                     mrb_mesh* out, mrb_qp_stats* stats) {
    MRB_CUDA(cudaSetDevice(ctx->device));
    memset(out, 0, sizeof(*out));
    const size_t n = (size_t)dims[0] * dims[1] * dims[2];
    int rc = MRB_OK, ns = 0;
    auto fail = [&](int code) { return code; };
    // intermediates live in the context's side buffers (grow-only, reused by the next call), not in the arena
    char* s0 = nullptr;
    const size_t gbytes = (n * 4 + 255) / 256 * 256;
    MRB_TRY(side_reserve(ctx, 0, 4 * gbytes, &s0));
    float *vol = (float*)s0, *smap = (float*)(s0 + gbytes), *fmap = (float*)(s0 + 2 * gbytes), *F = (float*)(s0 + 3 * gbytes);
    if ((rc = mrb_copy(ctx, vol, volume, n * 4)) != MRB_OK) return fail(rc);
    if ((rc = mrb_get_surface(ctx, vol, dims, nullptr, 0, &ns)) != MRB_OK) return fail(rc);
    char* s1 = nullptr;
    const size_t sbytes = ((size_t)(ns > 0 ? ns : 1) * 4 + 255) / 256 * 256;
    MRB_TRY(side_reserve(ctx, 1, 4 * sbytes, &s1));
    int* surf = (int*)s1;
    float* normals = (float*)(s1 + sbytes);
    if (ns > 0 && (rc = mrb_get_surface(ctx, vol, dims, surf, ns, &ns)) != MRB_OK) return fail(rc);
    if ((rc = mrb_index_map(ctx, dims, surf, ns, smap)) != MRB_OK) return fail(rc);
    if ((rc = mrb_surface_normals(ctx, vol, dims, surf, ns, normals)) != MRB_OK) return fail(rc);
    if ((rc = mrb_feature_map(ctx, vol, smap, normals, ns, dims, o.feature_threshold, fmap)) != MRB_OK) return fail(rc);
    if ((rc = mrb_optimize(ctx, vol, fmap, dims, 1, o.band_size, o.iters, F, stats)) != MRB_OK) return fail(rc);
    rc = mrb_mcubes_features(ctx, F, smap, normals, ns, dims, o.isolevel, o.feature_threshold, o.corner_threshold, device_out, out);
    if (rc == MRB_OK && device_out) {
        // the mesh lives in the arena (valid until the next call); nothing to do
    }
    return fail(rc);
}

int mrb_reconstruct(mrb_ctx* ctx, const float* volume, const int dims[3], const mrb_options* opt, int device_out,
                    mrb_mesh* out, mrb_qp_stats* stats) {
    MRB_REQUIRE(ctx && volume && dims && out, "NULL argument");
    mrb_options o;
    if (opt) o = *opt; else mrb_options_default(&o);
    MRB_REQUIRE(o.band_size >= 1 && o.band_size <= 64 && o.iters >= 0, "bad band_size / iters");
    if (o.using_features) return reconstruct_with_features(ctx, volume, dims, o, device_out, out, stats);
    MRB_CUDA(cudaSetDevice(ctx->device));
    MRB_TRY(ctx->arena.reset());
    memset(out, 0, sizeof(*out));
    Dims d = mkdims(dims);
    MRB_REQUIRE(d.count() <= (size_t)0x7fffffff, "grid too large");
    const float* dv = nullptr;
    MRB_TRY(stage_in(ctx, volume, d.count(), &dv));
    float* F = nullptr;
    MRB_TRY(arena_alloc(ctx, d.count(), &F));
    MRB_TRY(optimize_for_mc(ctx, dv, d, o.band_size, o.iters, o.isolevel, F, stats));
    MRB_CUDA(cudaEventRecord(ctx->ev[4], ctx->stream));
    float* verts = nullptr;
    int* faces = nullptr;
    int nv = 0, nf = 0;
    MRB_TRY(dev_mcubes(ctx, F, d, o.isolevel, &verts, &faces, &nv, &nf));
    MRB_CUDA(cudaEventRecord(ctx->ev[5], ctx->stream));
    MRB_TRY(export_mesh(ctx, verts, faces, nv, nf, device_out, out));
    if (stats) cudaEventElapsedTime(&stats->ms_mc, ctx->ev[4], ctx->ev[5]);
    return MRB_OK;
}

int mrb_reconstruct_batch(mrb_ctx* ctx, const float* volumes, const int dims[3], int nobj, const mrb_options* opt,
                          mrb_mesh* out, int* vert_off, int* face_off) {
    MRB_REQUIRE(ctx && volumes && dims && out && vert_off && face_off && nobj >= 0, "bad argument");
    mrb_options o;
    if (opt) o = *opt; else mrb_options_default(&o);
    MRB_REQUIRE(o.band_size >= 1 && o.band_size <= 64 && o.iters >= 0, "bad band_size / iters");
    MRB_REQUIRE(!o.using_features, "mrb_reconstruct_batch: the --feature-detection path runs one object at a time (mrb_reconstruct)");
    MRB_CUDA(cudaSetDevice(ctx->device));
    memset(out, 0, sizeof(*out));
    Dims d = mkdims(dims);
    MRB_REQUIRE(d.count() <= (size_t)0x7fffffff, "grid too large");
    HostMesh hm;
    vert_off[0] = 0; face_off[0] = 0;
    // Chunks of at most 2^28 voxels (1900 objects of 52^3; MRB_BATCH_CHUNK_LOG2 = 20..28 overrides): bounds the scratch and keeps
    // every index in 31 bits.  Chunks of 2^24 voxels keep the stage-3 working set in L2 (1024 x 52^3: stages 1-4 take 9 x 4.85 = 44
    // ms instead of 58 ms) but pay nine times the per-chunk round trips: 81 ms against 78 ms for the whole call, so one chunk it is.
    const bool stack = !getenv("MRB_BATCH_SEQUENTIAL");
    int chunk_log2 = 28;
    if (const char* e = getenv("MRB_BATCH_CHUNK_LOG2")) { chunk_log2 = atoi(e); if (chunk_log2 < 20) chunk_log2 = 20; if (chunk_log2 > 28) chunk_log2 = 28; }
    size_t per_chunk = ((size_t)1 << chunk_log2) / (d.count() ? d.count() : 1);
    if (per_chunk < 1) per_chunk = 1;
    for (int ob0 = 0; ob0 < nobj;) {
        const int ob1 = (int)std::min<size_t>((size_t)nobj, (size_t)ob0 + per_chunk);
        if (ob0 == 0) hm.grow_hint = (double)nobj / (double)(ob1 - ob0);
        int done = 0, rc = MRB_OK;
        if (stack) rc = batch_stacked(ctx, volumes, d, ob0, ob1, o, hm, vert_off, face_off, &done);
        if (rc == MRB_OK && !done) rc = batch_sequential(ctx, volumes, d, ob0, ob1, o, hm, vert_off, face_off);
        if (rc != MRB_OK) { hm.release(); return rc; }
        ob0 = ob1;
    }
    out->nverts = (int)(hm.nv3 / 3);
    out->nfaces = (int)(hm.nf3 / 3);
    out->on_device = 0;
    out->verts = hm.v ? hm.v : (float*)malloc(sizeof(float));
    out->faces = hm.f ? hm.f : (int*)malloc(sizeof(int));
    if (!out->verts || !out->faces) { set_error("mrb_reconstruct_batch: host malloc failed"); return MRB_ENOMEM; }
    return MRB_OK;
}

// ASCII PLY, mcubes.cpp:626-660.  "%g" == default ostream float formatting (6 significant digits).  The text is
// formatted by several host threads (one contiguous range of lines each) and written with one fwrite per range: the
// reference's `<< endl` per line makes this file write the slowest part of its whole run once the kernels are fast.
int mrb_write_ply(const char* path, const float* verts, int nverts, const int* faces, int nfaces, const float scale[3],
                  const float shift[3], int pad) {
    if (!path || (nverts > 0 && !verts) || (nfaces > 0 && !faces) || !scale || !shift) { set_error("mrb_write_ply: NULL argument"); return MRB_EINVAL; }
    FILE* f = fopen(path, "w");
    if (!f) { set_error("mrb_write_ply: cannot open %s", path); return MRB_EINVAL; }
    char head[512];
    int hl = snprintf(head, sizeof(head),
                      "ply\nformat ascii 1.0\nelement vertex %d\nproperty float x\nproperty float y\nproperty float z\n"
                      "element face %d\nproperty list uchar int vertex_indices\nend_header\n", nverts, nfaces);
    const size_t lines = (size_t)nverts + (size_t)nfaces;
    unsigned hw = std::thread::hardware_concurrency();
    size_t nth = hw ? hw : 4;
    if (nth > 32) nth = 32;
    if (lines < 20000) nth = 1;
    std::vector<std::string> part(nth);
    auto work = [&](size_t t) {
        const size_t l0 = lines * t / nth, l1 = lines * (t + 1) / nth;
        std::string& buf = part[t];
        buf.reserve((l1 - l0) * 28 + 64);
        char line[128];
        for (size_t l = l0; l < l1; l++) {
            int len;
            if (l < (size_t)nverts) {
                float p[3];
                // grid::getCloudPoint, grid.cpp:204-210
                for (int a = 0; a < 3; a++) p[a] = ((verts[3 * l + a] - (float)pad) * scale[a]) + shift[a];
                len = snprintf(line, sizeof(line), "%g %g %g\n", p[0], p[1], p[2]);
            } else {
                const size_t i = l - (size_t)nverts;
                len = snprintf(line, sizeof(line), "3 %d %d %d\n", faces[3 * i], faces[3 * i + 1], faces[3 * i + 2]);
            }
            buf.append(line, len);
        }
    };
    if (nth == 1) {
        work(0);
    } else {
        std::vector<std::thread> th;
        for (size_t t = 0; t < nth; t++) th.emplace_back(work, t);
        for (auto& t : th) t.join();
    }
    bool ok = fwrite(head, 1, hl, f) == (size_t)hl;
    for (size_t t = 0; t < nth && ok; t++) ok = fwrite(part[t].data(), 1, part[t].size(), f) == part[t].size();
    ok = (fclose(f) == 0) && ok;
    if (!ok) { set_error("mrb_write_ply: write to %s failed", path); return MRB_EINVAL; }
    return MRB_OK;
}

}  // extern "C"
