#!/bin/sh
# build_ref.sh -- TEST INFRASTRUCTURE.  Compiles the reference's OWN hot-path translation units, unmodified and where they
# lie under /root/reference, into oracle/_ref/ (git-ignored; nothing of the reference is copied into the repository):
#   grid/grid.cpp  distance_fields/dfields.cpp  narrow_band/narrowBand.cpp  optimize/primeqp.cpp  optimize/quadprog.cpp
#   marching_cubes/hasher.cpp  marching_cubes/mcubes.cpp            (+ oracle/ref_shim.cpp: flat-array C entry points)
# The reference's build system is not run (it needs Eigen, PCL and Boost, which are not installed): the third-party
# headers are replaced by the stand-ins of oracle/ref_stubs/ (see ref_eigen_standin.h for what that does not pin).
#   libmrb_ref.so      eigen-solver / SVD stand-in = include/mrb_eig3.h arithmetic (bit-comparable with the product)
#   libmrb_ref_front.so the front end (binvoxToPCL/, confidence_pcl/, voxelize/, grid.cpp) against PCL stand-ins, see below
#   libmrb_ref_dbl.so  eigen-solver / SVD stand-in = independent double-precision Jacobi (-DREF_STANDIN_DOUBLE)
# Flags follow the reference's CMake build (no -march, hence no FMA): -ffp-contract=off; gnu++03 like the source.
set -e
HERE=$(cd "$(dirname "$0")" && pwd)
REF=${MRB_REFERENCE_DIR:-/root/reference}
if [ ! -f "$REF/optimize/quadprog.cpp" ]; then
    echo "build_ref.sh: $REF not present (GPU box?) -- keeping whatever oracle/_ref/ holds" >&2
    exit 0
fi
CXX=/usr/bin/g++; [ -x "$CXX" ] || CXX=g++
OUT="$HERE/_ref"
mkdir -p "$OUT"
SRCS="$REF/grid/grid.cpp $REF/distance_fields/dfields.cpp $REF/narrow_band/narrowBand.cpp $REF/optimize/primeqp.cpp \
      $REF/optimize/quadprog.cpp $REF/marching_cubes/hasher.cpp $REF/marching_cubes/mcubes.cpp $HERE/ref_shim.cpp"
INC="-I$HERE/ref_stubs -I$HERE/../include -I$REF/grid -I$REF/distance_fields -I$REF/narrow_band -I$REF/optimize -I$REF/marching_cubes"
FLAGS="-std=gnu++03 -O2 -ffp-contract=off -fno-fast-math -fPIC -shared -w"
newer() {  # is $1 older than any source / stub?
    [ ! -f "$1" ] && return 0
    for f in $SRCS "$HERE"/ref_stubs/ref_eigen_standin.h "$HERE"/../include/mrb_eig3.h "$0"; do [ "$f" -nt "$1" ] && return 0; done
    return 1
}
if newer "$OUT/libmrb_ref.so"; then $CXX $FLAGS $INC $SRCS -o "$OUT/libmrb_ref.so"; fi
if newer "$OUT/libmrb_ref_dbl.so"; then $CXX $FLAGS -DREF_STANDIN_DOUBLE $INC $SRCS -o "$OUT/libmrb_ref_dbl.so"; fi
# The front end (row f1): binvoxToPcl.cpp, confidence_pcl/, voxelize.cpp, grid.cpp + oracle/ref_front_shim.cpp against the PCL
# stand-ins (exact-NN kd-tree, restated VoxelGrid, octree): libmrb_ref_front.so
FSRCS="$REF/grid/grid.cpp $REF/binvoxToPCL/binvoxToPcl.cpp $REF/voxelize/voxelize.cpp $REF/confidence_pcl/full_confidence.cpp \
       $REF/confidence_pcl/assign_confidence.cpp $REF/confidence_pcl/ConstConf.cpp $HERE/ref_front_shim.cpp"
FINC="-I$HERE/ref_stubs -I$HERE/../include -I$REF/grid -I$REF/binvoxToPCL -I$REF/voxelize -I$REF/confidence_pcl"
fnewer() {
    [ ! -f "$1" ] && return 0
    for f in $FSRCS "$HERE"/ref_stubs/pcl/filters/voxel_grid.h "$HERE"/ref_stubs/pcl/kdtree/kdtree_flann.h "$HERE"/ref_stubs/pcl/octree/octree.h "$0"; do
        [ "$f" -nt "$1" ] && return 0
    done
    return 1
}
if fnewer "$OUT/libmrb_ref_front.so"; then $CXX $FLAGS $FINC $FSRCS -o "$OUT/libmrb_ref_front.so"; fi
echo "oracle/_ref: $(ls "$OUT" | tr '\n' ' ')"
