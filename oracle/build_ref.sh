#!/usr/bin/env bash
# TEST INFRASTRUCTURE ONLY -- not part of the product path.
#
# Builds oracle/_ref/libcvref.so: the reference's OWN CPU implementation of the
# hot path (cv::warpAffine / cv::warpPerspective / cv::remap / cv::resize), compiled
# straight from the sources where they lie under /root/reference with plain g++.
# The reference's cmake build system is NOT run.  The handful of configuration
# headers cmake would normally emit (cvconfig.h, cv_cpu_config.h, custom_hal.hpp,
# opencv_modules.hpp, version_string.inc, *.simd_declarations.hpp) carry no
# algorithm; they are written below with the values of a default Linux x86-64
# configuration: baseline SSE3 (=> no FMA contraction in imgwarp.cpp/resize.cpp),
# pthreads parallel_for_ backend, no IPP, no OpenCL, no custom HAL, and the
# SSE4.1/AVX2 side files of imgproc (imgwarp.sse4_1.cpp, imgwarp.avx2.cpp,
# resize.sse4_1.cpp, resize.avx2.cpp) compiled with their ISA flags and selected at
# run time exactly as a stock build does.  core's *.dispatch.cpp files are built
# baseline-only (they are not on the hot path).
#
# Nothing from /root/reference is copied into the repository; outputs go only to
# oracle/_ref/ (git-ignored, shipped to the GPU box as a built artefact).
set -euo pipefail

HERE="$(cd "$(dirname "$0")" && pwd)"
REF="${OPENCV_REFERENCE_DIR:-/root/reference}"
OUT="$HERE/_ref"
OBJ="$OUT/obj"
GEN="$OUT/gen"
JOBS="${JOBS:-$(nproc)}"

if [ ! -d "$REF/modules/imgproc/src" ]; then
  echo "build_ref.sh: reference tree not found at $REF (expected on the GPU box: prebuilt _ref travels)" >&2
  exit 3
fi
if [ -f "$OUT/libcvref.so" ] && [ "$OUT/libcvref.so" -nt "$HERE/ref_shim.cpp" ] && [ "$OUT/libcvref.so" -nt "$0" ] && \
   [ -f "$OUT/libcvref_b200hal.so" ] && [ "$OUT/libcvref_b200hal.so" -nt "$HERE/../include/b200_geom.h" ] && \
   [ "$OUT/libcvref_b200hal.so" -nt "$HERE/../hal/cuda_b200/include/cuda_b200_hal.hpp" ]; then
  echo "build_ref.sh: $OUT/libcvref.so and libcvref_b200hal.so are up to date"
  exit 0
fi

mkdir -p "$OBJ" "$GEN/opencv2"

# ---- configuration headers (values of a default x86-64 Linux configuration) ----
cat > "$GEN/cvconfig.h" <<'EOF'
#ifndef OPENCV_CVCONFIG_H_INCLUDED
#define OPENCV_CVCONFIG_H_INCLUDED
#define BUILD_SHARED_LIBS
#define CV_ENABLE_INTRINSICS
#define CUDA_ARCH_BIN ""
#define CUDA_ARCH_FEATURES ""
#define CUDA_ARCH_PTX ""
#define HAVE_INTTYPES_H 1
#define HAVE_PTHREADS_PF
#define OPENCV_TRACE
#endif
EOF

cat > "$GEN/cv_cpu_config.h" <<'EOF'
// OpenCV CPU baseline features
#define CV_CPU_COMPILE_SSE 1
#define CV_CPU_BASELINE_COMPILE_SSE 1
#define CV_CPU_COMPILE_SSE2 1
#define CV_CPU_BASELINE_COMPILE_SSE2 1
#define CV_CPU_COMPILE_SSE3 1
#define CV_CPU_BASELINE_COMPILE_SSE3 1
#define CV_CPU_BASELINE_FEATURES 0 \
    , CV_CPU_SSE \
    , CV_CPU_SSE2 \
    , CV_CPU_SSE3 \

// OpenCV supported CPU dispatched features
#define CV_CPU_DISPATCH_COMPILE_SSE4_1 1
#define CV_CPU_DISPATCH_COMPILE_AVX2 1
#define CV_CPU_DISPATCH_FEATURES 0 \
    , CV_CPU_SSE4_1 \
    , CV_CPU_AVX2 \

EOF

cat > "$GEN/custom_hal.hpp" <<'EOF'
#ifndef _CUSTOM_HAL_INCLUDED_
#define _CUSTOM_HAL_INCLUDED_
#endif
EOF

cat > "$GEN/opencv2/opencv_modules.hpp" <<'EOF'
#define HAVE_OPENCV_CORE
#define HAVE_OPENCV_IMGPROC
EOF

cat > "$GEN/version_string.inc" <<'EOF'
"\n"
"General configuration for the hot-path oracle build (oracle/build_ref.sh)\n"
"  CPU/HW features: baseline SSE SSE2 SSE3; dispatched: SSE4_1 AVX2 (imgproc warp/resize side files)\n"
"  Parallel framework: pthreads\n"
"  IPP: NO   OpenCL: NO   Custom HAL: NO\n"
EOF

cat > "$GEN/opencv_data_config.hpp" <<'EOF2'
#define OPENCV_INSTALL_PREFIX "/usr/local"
#define OPENCV_DATA_INSTALL_PATH "share/opencv4"
#define OPENCV_BUILD_DIR "/nonexistent"
#define OPENCV_DATA_BUILD_DIR_SEARCH_PATHS ""
#define OPENCV_INSTALL_DATA_DIR_RELATIVE "../share/opencv4"
EOF2

# OpenCL is off: the generated kernel-source headers reduce to their include lines
for m in core imgproc stitching; do
cat > "$GEN/opencl_kernels_$m.hpp" <<'EOF2'
#include "opencv2/core/ocl.hpp"
#include "opencv2/core/ocl_genbase.hpp"
#include "opencv2/core/opencl/ocl_defs.hpp"
EOF2
done

# baseline-only dispatch declarations for every <name>.dispatch.cpp / <name>.simd.hpp we compile
gen_simd_decl() {  # $1 = module src dir, $2 = name
  local f="$GEN/$2.simd_declarations.hpp"
  {
    echo "#define CV_CPU_SIMD_FILENAME \"$1/$2.simd.hpp\""
    echo "#define CV_CPU_DISPATCH_MODES_ALL BASELINE"
    echo "#undef CV_CPU_SIMD_FILENAME"
  } > "$f"
}
for f in "$REF"/modules/core/src/*.simd.hpp; do
  gen_simd_decl "$REF/modules/core/src" "$(basename "$f" .simd.hpp)"
done
# calib3d's initUndistortRectifyMap is dispatched SSE2 / AVX2 in a stock build (modules/calib3d/CMakeLists.txt:
# ocv_add_dispatched_file(undistort SSE2 AVX2)); the AVX2 variant (4 fp64 lanes, FMA3) is what runs on any recent x86 host,
# and its vector body rounds differently from the scalar tail, so it is built here the way cmake would generate it.
cat > "$GEN/undistort.simd_declarations.hpp" <<EOF2
#define CV_CPU_SIMD_FILENAME "$REF/modules/calib3d/src/undistort.simd.hpp"
#define CV_CPU_DISPATCH_MODE AVX2
#include "opencv2/core/private/cv_cpu_include_simd_declarations.hpp"

#define CV_CPU_DISPATCH_MODES_ALL AVX2, BASELINE

#undef CV_CPU_SIMD_FILENAME
EOF2
cat > "$GEN/undistort.avx2.cpp" <<EOF2
#include "$REF/modules/calib3d/src/precomp.hpp"
#include "$REF/modules/calib3d/src/undistort.simd.hpp"
EOF2

INC=( -I"$GEN" -I"$REF/modules/core/include" -I"$REF/modules/imgproc/include" -I"$REF/modules/dnn/include" -I"$REF/modules/dnn/src"
      -I"$REF/modules/calib3d/include" -I"$REF/modules/calib3d/src" -I"$REF/modules/features2d/include" -I"$REF/modules/flann/include"
      -I"$REF/modules/core/src" -I"$REF/modules/imgproc/src"
      -I"$REF/modules/stitching/include" -I"$REF/modules/stitching/src"
      -I"$REF/3rdparty/include" )
CXXFLAGS=( -std=c++17 -O3 -DNDEBUG -fPIC -fvisibility=default -msse3 -ffp-contract=off
           -fsigned-char -pthread -D__OPENCV_BUILD=1 -D_USE_MATH_DEFINES
           -D__STDC_CONSTANT_MACROS -D__STDC_LIMIT_MACROS -D__STDC_FORMAT_MACROS
           -w )

# ---- source list: all of core (runtime the path depends on) + the hot-path files of imgproc ----
SRCS=()
for f in "$REF"/modules/core/src/*.cpp "$REF"/modules/core/src/parallel/*.cpp "$REF"/modules/core/src/utils/*.cpp; do
  case "$(basename "$f")" in
    # device / platform back-ends that are not part of a default CPU-only build
    cuda_*.cpp|directx.cpp|opengl.cpp|gl_core_3_1.cpp|ovx.cpp|va_intel.cpp|parallel_openmp.cpp|parallel_tbb.cpp) continue;;
  esac
  SRCS+=( "$f" )
done
IMG="$REF/modules/imgproc/src"
SRCS+=( "$IMG/imgwarp.cpp" "$IMG/resize.cpp" "$IMG/tables.cpp" "$IMG/pyramids.cpp" )
# dnn::blobFromImages (a caller of cv::resize the library fuses): this one file compiles on its own
SRCS+=( "$REF/modules/dnn/src/dnn_utils.cpp" )
# cv::fisheye::initUndistortRectifyMap (the map generator of BASELINE config 4): fisheye.cpp compiles on its own; the three
# calib3d functions its OTHER entry points call (solvePnP, solvePnPRansac, Rodrigues) are stubbed in ref_shim.cpp
SRCS+=( "$REF/modules/calib3d/src/fisheye.cpp" "$REF/modules/calib3d/src/undistort.dispatch.cpp" )
# stitching: the rotation warpers (map builders + cv::remap); the rest of the module is not needed
SRCS+=( "$REF/modules/stitching/src/warpers.cpp" )
SIDE_SSE41=( "$IMG/imgwarp.sse4_1.cpp" "$IMG/resize.sse4_1.cpp" )
SIDE_AVX2=( "$IMG/imgwarp.avx2.cpp" "$IMG/resize.avx2.cpp" )
DISP_AVX2=( "$GEN/undistort.avx2.cpp" )       # generated dispatch variants: also need CV_CPU_DISPATCH_MODE

# ---- emit a throwaway Makefile so the ~110 TUs build in parallel ----
MK="$OUT/build.mk"
{
  echo "all: $OUT/libcvref.so"
  OBJS=()
  emit() { # src extra-flags
    local src="$1"; shift
    local o="$OBJ/$(echo "${src#$REF/modules/}" | tr '/' '_' | sed 's/\.cpp$/.o/')"
    OBJS+=( "$o" )
    echo "$o: $src"
    echo "	@echo CXX ${src#$REF/}"
    echo "	@g++ ${CXXFLAGS[*]} $* ${INC[*]} -c $src -o $o"
  }
  for s in "${SRCS[@]}"; do emit "$s"; done
  for s in "${SIDE_SSE41[@]}"; do emit "$s" -mssse3 -msse4.1 -DCV_CPU_COMPILE_SSSE3=1 -DCV_CPU_COMPILE_SSE4_1=1; done
  for s in "${SIDE_AVX2[@]}"; do emit "$s" -mssse3 -msse4.1 -mpopcnt -msse4.2 -mavx -mf16c -mavx2 -mfma -DCV_CPU_COMPILE_SSSE3=1 -DCV_CPU_COMPILE_SSE4_1=1 -DCV_CPU_COMPILE_POPCNT=1 -DCV_CPU_COMPILE_SSE4_2=1 -DCV_CPU_COMPILE_AVX=1 -DCV_CPU_COMPILE_FP16=1 -DCV_CPU_COMPILE_AVX2=1 -DCV_CPU_COMPILE_FMA3=1; done
  for s in "${DISP_AVX2[@]}"; do emit "$s" -mssse3 -msse4.1 -mpopcnt -msse4.2 -mavx -mf16c -mavx2 -mfma -DCV_CPU_COMPILE_SSSE3=1 -DCV_CPU_COMPILE_SSE4_1=1 -DCV_CPU_COMPILE_POPCNT=1 -DCV_CPU_COMPILE_SSE4_2=1 -DCV_CPU_COMPILE_AVX=1 -DCV_CPU_COMPILE_FP16=1 -DCV_CPU_COMPILE_AVX2=1 -DCV_CPU_COMPILE_FMA3=1 -DCV_CPU_DISPATCH_MODE=AVX2; done
  SH="$OBJ/ref_shim.o"
  echo "$SH: $HERE/ref_shim.cpp"
  echo "	@g++ ${CXXFLAGS[*]} ${INC[*]} -c $HERE/ref_shim.cpp -o $SH"
  echo "$OUT/libcvref.so: ${OBJS[*]} $SH"
  echo "	@echo LINK libcvref.so"
  echo "	@g++ -shared -o $OUT/libcvref.so ${OBJS[*]} $SH -pthread -ldl -lm"
} > "$MK"

make -f "$MK" -j"$JOBS" all
echo "build_ref.sh: built $OUT/libcvref.so"

# ---- second flavour: the SAME reference sources with hal/cuda_b200 registered ------------------------------
# libcvref_b200hal.so = reference cv::warpAffine / warpPerspective / remap / resize whose CALL_HAL sites
# (imgwarp.cpp:1608-1622,:2255,:2749; resize.cpp:3648) reach libb200geom.so through hal/cuda_b200's header.
# Only custom_hal.hpp differs; core objects are shared.  Used by tests/test_gpu_hal_dropin.py on the GPU box.
B200LIB="$HERE/../opencv_b200/lib/libb200geom.so"
if [ -f "$B200LIB" ]; then
  GENH="$OUT/gen_hal"; OBJH="$OUT/obj_hal"; mkdir -p "$GENH" "$OBJH"
  cat > "$GENH/custom_hal.hpp" <<'EOF2'
#ifndef _CUSTOM_HAL_INCLUDED_
#define _CUSTOM_HAL_INCLUDED_
#include "cuda_b200_hal.hpp"
#endif
EOF2
  HALINC=( -I"$GENH" -I"$HERE/../hal/cuda_b200/include" -I"$HERE/../include" )
  HOBJS=()
  for src in "$IMG/imgwarp.cpp" "$IMG/resize.cpp" "$IMG/pyramids.cpp" "$HERE/ref_shim.cpp"; do
    o="$OBJH/$(basename "$src" .cpp).o"; HOBJS+=( "$o" )
    if [ ! -f "$o" ] || [ "$src" -nt "$o" ] || [ "$HERE/../hal/cuda_b200/include/cuda_b200_hal.hpp" -nt "$o" ] || [ "$HERE/../include/b200_geom.h" -nt "$o" ]; then
      echo "CXX(hal) $(basename "$src")"
      g++ "${CXXFLAGS[@]}" "${HALINC[@]}" "${INC[@]}" -c "$src" -o "$o" &
    fi
  done
  wait
  # every object of the plain build except the three recompiled ones
  REST=$(ls "$OBJ"/*.o | grep -v -e imgproc_src_imgwarp.o -e imgproc_src_resize.o -e imgproc_src_pyramids.o -e ref_shim.o)
  g++ -shared -o "$OUT/libcvref_b200hal.so" $REST "${HOBJS[@]}" -L"$(dirname "$B200LIB")" -lb200geom \
      -Wl,-rpath,'$ORIGIN/../../opencv_b200/lib' -pthread -ldl -lm
  echo "build_ref.sh: built $OUT/libcvref_b200hal.so (reference + hal/cuda_b200)"
fi
