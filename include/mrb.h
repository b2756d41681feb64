/*
 * mrb.h -- C-ABI of mesh_reconstruction_b200 (libmrb.so).
 *
 * Drop-in boundary for the volumetric surface-extraction hot path of
 * CRLab/Mesh_Reconstruction:  occupancy grid -> exact EDT -> narrow band ->
 * bound-constrained smoothness QP (damped projected Jacobi) -> marching cubes
 * with hash vertex de-duplication.  The reference has no FFI layer; its
 * boundary is the C++ module API called from mesh_reconstruction.cpp:119-134.
 * Every entry point below names the reference function it replaces.  The C++
 * shim in compat/ re-creates the reference signatures (gridPtr etc.) on top of
 * these symbols; INTEGRATION.md shows the binding.
 *
 * Conventions
 *  - Grids are flat fp32 arrays in the reference's storage order
 *    voxels[x][y][z] (z contiguous, grid.cpp:41-48): element (x,y,z) lives at
 *    (x*dims[1] + y)*dims[2] + z.  dims = {nx, ny, nz}.
 *  - "Linear indexes" handed across the API (findIndexes, getSurface) are the
 *    reference's x-fastest index  z*nx*ny + y*nx + x  (grid.cpp:172-174).
 *  - Every pointer argument may be a host pointer or a device pointer
 *    (detected with cudaPointerGetAttributes); host buffers are staged through
 *    the context's stream.  No entry point falls back to a CPU implementation:
 *    without a usable GPU mrb_ctx_create fails.
 *  - All functions return 0 on success, a negative MRB_E* code on failure;
 *    mrb_last_error() returns the message of the calling thread's last failure.
 *    (The reference prints and exit(1)s -- the compat shim reproduces that.)
 *  - One mrb_ctx per host thread / per GPU; calls on one ctx are serialised on
 *    its stream and are synchronous on return unless stated otherwise.
 *  - Dead-man switch: the persistent sweep kernels (stage 3, single- and multi-GPU) and the flipEdges
 *    kernel wait on progress flags of other blocks / peer GPUs in spin loops that execute a trap after
 *    about 2^26 polls (a lost peer or predecessor).  The call then fails with MRB_ECUDA and the error is
 *    STICKY: the CUDA context of the whole process is unusable afterwards (every mrb_ctx and every other
 *    CUDA library in it); the process has to be restarted or its primary context reset with cudaDeviceReset.
 */
#ifndef MRB_H
#define MRB_H

#include <stdint.h>
#include <stddef.h>

#ifdef __cplusplus
extern "C" {
#endif

#define MRB_OK            0
#define MRB_EINVAL       -1   /* bad argument */
#define MRB_ECUDA        -2   /* CUDA runtime error */
#define MRB_ENOMEM       -3   /* allocation failed */
#define MRB_EBORDER      -4   /* band touches the grid border (reference: exit(1), grid.cpp:20-37) */
#define MRB_ECAPACITY    -5   /* caller-provided buffer too small */
#define MRB_EHASH        -6   /* vertex hash lookup failed (reference: mcubes.cpp:652-656) */

typedef struct mrb_ctx mrb_ctx;

/* Reference constants (quadprog.cu:237, primeqp.cpp:219, dfields.cpp:50). */
#define MRB_BAND_SIZE_DEFAULT   4
#define MRB_QP_ITERS_DEFAULT    500
#define MRB_SQDIST_CLAMP        100000

/* ---- context ---------------------------------------------------------- */
/* Creates a context on CUDA device `device` (stream + grow-only scratch arena). */
int  mrb_ctx_create(int device, mrb_ctx** out);
void mrb_ctx_destroy(mrb_ctx* ctx);
/* The CUDA stream (cudaStream_t) all work of this ctx is issued on. */
void* mrb_ctx_stream(mrb_ctx* ctx);
const char* mrb_last_error(void);
/* Library build info: "mrb <version> sm_100a". */
const char* mrb_version(void);
/* Number of kernel launches issued by this ctx since creation (bench: gpu_launches). */
uint64_t mrb_ctx_launch_count(mrb_ctx* ctx);
/* Utility: copies nbytes between any two host/device buffers on the ctx stream and waits
 * (used by bindings to take device-resident results that live in the ctx arena). */
int  mrb_copy(mrb_ctx* ctx, void* dst, const void* src, size_t nbytes);

/* ---- stage 1: distance field (distance_fields/) ------------------------ */
/* fastPerim (dfields.cpp:7-37) / fastPerimGPU (dfields.cu:62-89):
 * out = 1, or 0 if any voxel of the border-clamped 3x3x3 window differs from the centre. */
int mrb_fast_perim(mrb_ctx* ctx, const float* volume, const int dims[3], float* out);
/* getsqdist (dfields.cpp:41-99) / sqdistpt1-3 (dfields.cu:91-151):
 * out = min(100000, exact squared Euclidean distance to the nearest voxel with in == 0). */
int mrb_sqdist(mrb_ctx* ctx, const float* in, const int dims[3], float* out);
/* getsqrt (dfields.cpp:182-192) / sqrtgpu (dfields.cu:153-159): correctly rounded sqrtf. */
int mrb_sqrt(mrb_ctx* ctx, const float* in, const int dims[3], float* out);
/* dfield_gpu (dfields.cu:161-200) == getsqrt(getsqdist(fastPerim(volume))). */
int mrb_dfield(mrb_ctx* ctx, const float* volume, const int dims[3], float* margin);

/* ---- feature pre-pass (distance_fields/, only consumed with --feature-detection) --------- */
/* The normals come from Eigen::SelfAdjointEigenSolver in the reference; here from the deterministic
 * Jacobi solver of include/mrb_eig3.h (parity unpinned against the reference, exact against our oracle). */
/* getSurface (dfields.cpp:322-338): x-fastest linear indexes of occupied voxels (== 1) with an empty (== 0)
 * voxel in their 3x3x3 window, in (x,y,z) order.  idx_out may be NULL (count only). */
int mrb_get_surface(mrb_ctx* ctx, const float* volume, const int dims[3], int* idx_out, int capacity, int* n_out);
/* getSurfaceNormals (dfields.cpp:341-349): one unit normal (3 floats) per surface voxel. */
int mrb_surface_normals(mrb_ctx* ctx, const float* volume, const int dims[3], const int* surface, int n,
                        float* normals_out);
/* getFeatureMap (dfields.cpp:447-555): 1 on the surface voxels of every boundary cell whose minimum pairwise
 * normal dot product is below `threshold`, else 0.  surface_map = getIndexMap(volume, surface). */
int mrb_feature_map(mrb_ctx* ctx, const float* volume, const float* surface_map, const float* normals, int nnormals,
                    const int dims[3], float threshold, float* feature_map_out);

/* ---- stage 2: narrow band (narrow_band/, grid/) ------------------------ */
/* erode_grid / dilate_grid (narrowBand.cpp:10-79): 6-neighbour, interior voxels only. */
int mrb_erode(mrb_ctx* ctx, const float* in, const int dims[3], float* out);
int mrb_dilate(mrb_ctx* ctx, const float* in, const int dims[3], float* out);
/* createBands (narrowBand.cpp:82-102): band0 = margin<=band_size; tight = erode(band0);
 * band = dilate(tight).  band/tight are 0/1 float grids. */
int mrb_create_bands(mrb_ctx* ctx, const float* margin, const int dims[3], float band_size,
                     float* band, float* tight);
/* findIndexes (grid.cpp:350-371): non-zero voxels in (x,y,z)-lexicographic order, values are
 * x-fastest linear indexes.  idx_out may be NULL (count only); otherwise capacity entries. */
int mrb_find_indexes(mrb_ctx* ctx, const float* grid, const int dims[3],
                     int* idx_out, int capacity, int* n_out);
/* getIndexMap (grid.cpp:374-389): map = -1 everywhere, map(indexes[i]) = (float)i. */
int mrb_index_map(mrb_ctx* ctx, const int dims[3], const int* indexes, int n, float* map_out);

/* mrb_find_indexes for callers that must not allocate.  No counterpart in the reference.
 *
 * Same result as mrb_find_indexes (band voxels in reference order, x-fastest linear indexes), but
 * the call makes no cudaMalloc / cudaFree / allocating library call of any kind and needs no
 * mrb_ctx.  It exists for programs that cannot let a library allocate device memory: the card's
 * memory already belongs to the host application's own pool (a deep-learning framework's caching
 * allocator, a fixed perception arena on a robot), or the call sits in a loop with a deadline
 * where the device-wide stall of cudaMalloc/cudaFree is not acceptable, or the process runs under
 * a fixed memory quota.  Its scratch (one counter per 1024-voxel tile) is a fixed 64.25 KiB array
 * inside the library image, held by every GPU that loads libmrb until the process exits.
 *
 *  - grid_dev, idx_out_dev: DEVICE pointers provided by the caller (host buffers are rejected:
 *    staging them would need an allocation).  idx_out_dev may be NULL to get the count only.
 *  - *n_out (host) receives the number of non-zero voxels; capacity = entries idx_out_dev can hold.
 *  - Limit: dims[0]*dims[1]*dims[2] <= MRB_NOALLOC_MAX_VOXELS (256^3); larger grids are refused
 *    with MRB_ECAPACITY and a message -- use mrb_find_indexes for them.
 *  - stream: a cudaStream_t (NULL = the default stream); the call returns after the work finished.
 *  - The first call on each GPU looks up the scratch array's device address (the current device
 *    must be the one the buffers live on).  Make that first call before starting any stream capture.
 *  - All calls on a GPU share the one array, which keeps the previous call's contents (the part a
 *    call reads is cleared first).  Two calls that overlap on the same GPU -- different streams or
 *    different host threads -- race on it: serialise them. */
#define MRB_NOALLOC_MAX_VOXELS  (256 * 256 * 256)
int mrb_find_indexes_noalloc(const float* grid_dev, const int dims[3], int* idx_out_dev, int capacity,
                             int* n_out, void* stream);

/* ---- stage 3: QP (optimize/) ------------------------------------------- */
/* primeQP (primeqp.cpp:168-223) + getIr/getJc/getPr (quadprog.cpp:8-42).
 * n must equal the band count.  Outputs (each may be NULL to skip):
 *   lb, ub, x0, invdg : n floats;  jc : n+1 ints;  ir, pr : nnz_capacity entries;
 *   *nnz_out = number of off-diagonal non-zeros of R = H - diag(H).
 * Fails with MRB_EBORDER when a tight-band voxel lies on the grid border. */
int mrb_prime_qp(mrb_ctx* ctx, const float* volume, const float* margin,
                 const float* band, const float* tight, const int dims[3], int n,
                 float* lb, float* ub, float* x0, float* invdg,
                 int* jc, int* ir, float* pr, int nnz_capacity, int* nnz_out);
/* runQP (quadprog.cpp:64-105) / runQPGPU (quadprog.cu:107-186): `iters` damped projected
 * Jacobi sweeps on an explicit CSC matrix, then the reference's feature reset loop. */
int mrb_run_qp(mrb_ctx* ctx, int n, const int* jc, const int* ir, const float* pr,
               const float* invdg, const float* lb, const float* ub, const float* x0,
               int iters, const int* feature_idx, int nfeat, int using_features,
               float* x_out);

typedef struct mrb_qp_stats {
    int     n_band;          /* B: band voxels */
    int     n_tight;         /* T: tight-band voxels */
    int     iters;           /* sweeps executed */
    float   last_max_delta;  /* max |x_new - x_old| over the band in the last sweep (warp-shuffle reduced) */
    float   ms_dfield, ms_band, ms_qp, ms_mc;   /* device time per stage of the last fused call */
    float   ms_qp_kernel;    /* device time of the persistent sweep kernel alone (roofline denominator) */
} mrb_qp_stats;

/* optimize (quadprog.cu:236-280): stages 1-3 device-resident; F = +-(band_size+1) off band,
 * QP solution on the band.  feature_map may be NULL when !using_features.
 * band_size / iters: reference constants are 4 / 500 (hard-coded there). */
int mrb_optimize(mrb_ctx* ctx, const float* volume, const float* feature_map, const int dims[3],
                 int using_features, int band_size, int iters, float* F_out, mrb_qp_stats* stats);

/* ---- stage 4: marching cubes + hasher (marching_cubes/) ----------------- */
typedef struct mrb_mesh {
    float*  verts;     /* nverts * 3, grid coordinates (before getCloudPoint), first-appearance order */
    int*    faces;     /* nfaces * 3 vertex ids, reference triangle order */
    int     nverts;
    int     nfaces;
    int     on_device; /* where verts/faces live = the device_out the call was made with:
                        * 0 malloc'd host arrays owned by the caller (mrb_mesh_free);
                        * 1 device memory of the context's arena, 2 pinned host memory of the context (no extra host copy):
                        *   both owned by the context and valid until the next call on it; mrb_mesh_free only clears the struct */
} mrb_mesh;

/* mcubes (mcubes.cpp:592-662) without the PLY write, default (non-feature) path:
 * getGridCells + Polygonise + pointHashmap.  device_out: see mrb_mesh.on_device (0, 1 or 2). */
int  mrb_mcubes(mrb_ctx* ctx, const float* F, const int dims[3], float isolevel,
                int device_out, mrb_mesh* out);
void mrb_mesh_free(mrb_mesh* mesh);
/* mcubes with USING_FEATURES (mcubes.cpp:592-662): isFeature per cell, least-squares feature points with triangle
 * fans, flipEdges between neighbouring feature cells, then the same hash de-duplication.  surface_map / normals are
 * getIndexMap(volume, getSurface(volume)) and getSurfaceNormals.  The feature point solve is our own definition
 * (include/mrb_eig3.h, include/mrb_feature_mc.h) in place of Eigen::JacobiSVD: parity unpinned against the
 * reference, bit-exact against our oracle. */
int  mrb_mcubes_features(mrb_ctx* ctx, const float* F, const float* surface_map, const float* normals, int nnormals,
                         const int dims[3], float isolevel, float feat_thresh, float corner_thresh,
                         int device_out, mrb_mesh* out);
/* hashXYZ (hasher.h:99-104) for n points (xyz interleaved); keys_out: n uint64. */
int  mrb_hash_points(mrb_ctx* ctx, const float* xyz, int n, const int dims[3], uint64_t* keys_out);
/* ASCII PLY exactly as mcubes.cpp:626-660 (vertices through grid::getCloudPoint, grid.cpp:204-210). */
int  mrb_write_ply(const char* path, const float* verts, int nverts, const int* faces, int nfaces,
                   const float scale[3], const float shift[3], int pad);

/* ---- fused path --------------------------------------------------------- */
typedef struct mrb_options {
    int   using_features;     /* --feature-detection (0: default path) */
    int   band_size;          /* 4 */
    int   iters;              /* 500 */
    float isolevel;           /* 0.0 (mesh_reconstruction.cpp:134) */
    float feature_threshold;  /* 0.75 */
    float corner_threshold;   /* 0.8 */
} mrb_options;
void mrb_options_default(mrb_options* o);
/* optimize + mcubes with everything resident on the device between the stages. */
int mrb_reconstruct(mrb_ctx* ctx, const float* volume, const int dims[3], const mrb_options* opt,
                    int device_out, mrb_mesh* out, mrb_qp_stats* stats);

/* Object-parallel batch (BASELINE config 5: many independent small objects): the fused path on `nobj`
 * grids of identical dims stored back to back ([nobj][nx][ny][nz], host or device memory).  The meshes are
 * returned concatenated (host memory, release with mrb_mesh_free): object o owns vertices
 * [vert_off[o], vert_off[o+1]) and faces [face_off[o], face_off[o+1]); face entries are vertex ids local
 * to their object, exactly what mcubes (mcubes.cpp:592-662) would write for that object alone.
 * vert_off / face_off: nobj+1 ints each (host). */
int mrb_reconstruct_batch(mrb_ctx* ctx, const float* volumes, const int dims[3], int nobj,
                          const mrb_options* opt, mrb_mesh* out, int* vert_off, int* face_off);

/* ---- stepwise stage 3 for slab-sharded multi-GPU runs ---------------------------------------
 * No counterpart in the reference (it is single-GPU).  A plan holds stages 1-2 and the 2x2x4 bricks of one binary
 * occupancy grid; bricks are numbered with x slowest, so an x-slab is a contiguous brick range and its halo is
 * the neighbouring brick plane.  The iterate lives in caller-owned device buffers of 16*(Q+1) floats (four SoA
 * planes of Q+1 float4; slot Q of every plane must stay 0).  One sweep over [q_begin, q_end) is exactly one
 * doStep (quadprog.cpp:45-61) for every voxel of those bricks; running all bricks for `iters` sweeps with
 * alternating buffers and assembling gives the same F as mrb_optimize, bit for bit. */
typedef struct mrb_qp_plan mrb_qp_plan;
int  mrb_qp_plan_create(mrb_ctx* ctx, const float* volume, const int dims[3], int band_size, mrb_qp_plan** out);
void mrb_qp_plan_destroy(mrb_qp_plan* plan);
int  mrb_qp_plan_info(mrb_qp_plan* plan, int* n_bricks, int* n_band, int* n_tight, int* n_brick_planes);
/* off[X] = first brick id of brick plane X (voxel planes 2X, 2X+1); n_brick_planes + 1 entries. */
int  mrb_qp_plan_plane_offsets(mrb_qp_plan* plan, int* off);
int  mrb_qp_plan_init_x(mrb_qp_plan* plan, float* x_dev);
int  mrb_qp_plan_sweep(mrb_qp_plan* plan, const float* x_in_dev, float* x_out_dev, int q_begin, int q_end);
int  mrb_qp_plan_assemble(mrb_qp_plan* plan, const float* x_dev, float* F_out);
/* Slab-sharded solve in ONE persistent kernel per GPU: rank `rank` of `world` sweeps the bricks of its brick planes
 * [plane_cuts[rank], plane_cuts[rank+1]) for `iters` sweeps and stores the bricks of its first / last plane
 * directly into the neighbour ranks' iterate buffers over NVLink; per-sweep flags in the neighbours' memory replace
 * a host-driven exchange.  xa_ptrs / xb_ptrs / flag_ptrs: `world` peer-mapped device pointers each, e.g. from torch
 * symmetric memory: every rank's buffer A and buffer B (16 * (Q + 9) floats each, 128-byte aligned; during the solve
 * they hold a slab-local layout) and a block of 512 uint32 that the caller zeroes before a barrier across the ranks.
 * All ranks must call this concurrently with the same plan contents.  On return the owned brick range of buffer A
 * (iters even) or B (iters odd) holds the result in the plan's layout (four planes of Q + 1 float4). */
int  mrb_qp_plan_solve_sharded(mrb_qp_plan* plan, int rank, int world, void* const* xa_ptrs, void* const* xb_ptrs,
                               void* const* flag_ptrs, const int* plane_cuts, int iters);
float mrb_qp_plan_last_kernel_ms(mrb_qp_plan* plan);
/* Slab-local plans (stages 1-2 sharded too).  mrb_qp_plan_create_slab builds a plan of a SUB-grid, planes [a, b) of a larger grid
 * whose x faces are cuts: everything at least band_size + 5 voxel planes away from a cut equals what the whole grid's plan holds
 * (perimeter +-1, EDT +-(band_size+1), erosion +-1, dilation +-1, the brick build's neighbour tests +-1), so a rank takes its
 * own planes plus that margin.  A band that touches a face is not an error here (it always touches the cuts); run the unsharded
 * path to get MRB_EBORDER.  mrb_qp_plan_solve_sharded_sub is mrb_qp_plan_solve_sharded on such a plan: the rank owns brick planes
 * [plane_lo, plane_hi) of ITS plan, which must also hold the neighbours' adjacent planes; nb_layout = {owned bricks, lower-halo
 * bricks, upper-halo bricks} of the lower neighbour, then of the upper neighbour (exchanged by the caller; ignored where there is no
 * neighbour).  Result as in mrb_qp_plan_solve_sharded, in this plan's layout.  No counterpart in the reference. */
int  mrb_qp_plan_create_slab(mrb_ctx* ctx, const float* volume_sub, const int dims_sub[3], int band_size, mrb_qp_plan** out);
int  mrb_qp_plan_solve_sharded_sub(mrb_qp_plan* plan, int rank, int world, void* const* xa_ptrs, void* const* xb_ptrs,
                                   void* const* flag_ptrs, int plane_lo, int plane_hi, const int* nb_layout, int iters);
/* mrb_qp_plan_assemble for the voxel planes [x_begin, x_end) only: F_slab_out holds (x_end - x_begin) * ny * nz floats.  Only
 * the bricks of those planes are read from x_dev (plan layout), so a rank of a slab-sharded run needs its own brick planes and
 * the first plane of its upper neighbour.  No counterpart in the reference. */
int  mrb_qp_plan_assemble_slab(mrb_qp_plan* plan, const float* x_dev, int x_begin, int x_end, float* F_slab_out);
/* mrb_mcubes on an x-slab of a larger grid: F_slab = planes [x_origin, x_origin + dims[0]) of the whole grid (dims[1], dims[2] are
 * the whole grid's).  Cells x_origin .. x_origin + dims[0] - 2; vertex coordinates and hash keys are those of the whole grid, bit for
 * bit; faces index this slab's own vertex list.  Concatenating the slabs' triangles in slab order and merging the vertices of the
 * shared planes by hash key (first slab wins) gives mcubes' result on the whole grid (mesh_reconstruction_b200/slab.py:
 * merge_slab_meshes).  No counterpart in the reference (marching_cubes/mcubes.cpp:592-662 is single-process). */
int  mrb_mcubes_slab(mrb_ctx* ctx, const float* F_slab, const int dims[3], int x_origin, float isolevel, int device_out,
                     mrb_mesh* out);

/* ---- front end: binvox completion + observed points -> occupancy grid -------------------------
 * The reference does this on the CPU with PCL (mesh_reconstruction.cpp:99-120).  Points are xyz interleaved floats, finite.
 * Outputs follow mrb_mesh's device_out convention (0 malloc'd host array, 1 the context's arena, 2 its pinned staging; 1 and 2
 * are valid until the next call on the context). */
typedef struct mrb_points { float* xyz; int n; int on_device; } mrb_points;
typedef struct mrb_volume {
    float* voxels;          /* dims[0]*dims[1]*dims[2] floats, 0 / 1, [x][y][z] */
    int    dims[3];
    float  scale[3], shift[3];   /* grid::scale / grid::shift (grid.cpp:94-120): voxel -> world by getCloudPoint */
    int    pad;             /* 6 empty voxels on every side (grid.cpp:91) */
    int    occupied_leaves; /* "PointCloud after filtering" (voxelize.cpp:48) */
    int    on_device;
} mrb_volume;
typedef struct mrb_front_info {
    float observed_dist_total;  int observed_count;     /* getDensity(observed): "Total Dist: .. Count: .." (binvoxToPcl.cpp:196) */
    float completed_dist_total; int completed_count;    /* getDensity(binvoxToPCL(file, 1)) */
    float observed_density, completed_density;          /* binvoxToPcl.cpp:205-209 (observed already scaled by 3/8) */
    int   res_factor;
    int   n_predicted;          /* points of binvoxToPCL(file, res_factor) */
    int   n_cloud;              /* observed + predicted = "PointCloud before filtering" */
    float leaf;                 /* getResolution: "Leaf size is" (voxelize.cpp:92-93) */
} mrb_front_info;
/* binvoxToPCL (binvoxToPCL/binvoxToPcl.cpp:105-169) on the decoded voxel bytes of a binvox file (read_binvox, :15-103, stays host
 * I/O): dwh = {depth, width, height}, xform = {scale, tx, ty, tz}; voxel (x, y, z) is voxels[x*width*height + z*width + y]
 * (get_index, :9-13).  Points in the reference's order (cell x, y, z, then sample i, j, k).  res_factor < 1 is MRB_EINVAL
 * (the reference exits). */
int  mrb_binvox_to_points(mrb_ctx* ctx, const unsigned char* voxels, const int dwh[3], const float xform[4], int res_factor,
                          int device_out, mrb_points* out);
void mrb_points_free(mrb_points* p);
/* The statistic inside getDensity (binvoxToPcl.cpp:172-201) and getResolution (voxelize/voxelize.cpp:6-32): for every tenth point
 * the distance to its nearest neighbour among the other nine tenths, summed in float in query order.  density = count / dist_total,
 * leaf size = 2 * dist_total / count.  pcl::KdTreeFLANN itself is not in the reference tree: exact search, squared distance
 * ((dx^2 + dy^2) + dz^2) in float. */
int  mrb_nn_stats(mrb_ctx* ctx, const float* xyz, int n, float* dist_total, int* count);
/* voxelize (voxelize.cpp:34-51) + grid(cloud, vox) (grid/grid.cpp:84-121) + fillGrid(res_factor) (grid.cpp:213-265; skipped for
 * res_factor <= 0, as m_cubes does) + getBinaryVolume (grid.cpp:322-334) with every strength 1 (ConstConf(1),
 * mesh_reconstruction.cpp:107).  pcl::VoxelGrid is restated: leaf (i, j, k) = floor(p * (1 / leaf)) - floor(min * (1 / leaf)),
 * centroid = float sum of the leaf's points in cloud order / count. */
int  mrb_voxelize_leaf(mrb_ctx* ctx, const float* xyz, int n, float leaf, int res_factor, int device_out, mrb_volume* out);
void mrb_volume_free(mrb_volume* v);
/* mesh_reconstruction.cpp:99-120 in one call: getResolutionFactor (res_factor < 1; a value >= 1 is taken as given), binvoxToPCL,
 * observed points followed by the completion, voxelizeData, createGrid, getBinaryVolume.  With device_out = 1 the volume can be
 * handed straight to mrb_reconstruct (which stages its own copy before the arena is reused).  info may be NULL. */
int  mrb_front_end(mrb_ctx* ctx, const unsigned char* voxels, const int dwh[3], const float xform[4], const float* observed,
                   int n_observed, int res_factor, int device_out, mrb_volume* out, mrb_front_info* info);

#ifdef __cplusplus
}
#endif
#endif /* MRB_H */
