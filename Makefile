# Builds libb200geom.so (sm_100a only) and the test oracle.
NVCC ?= nvcc
ARCH := -gencode arch=compute_100a,code=sm_100a
# -fmad=false: the reference's default build has no FMA; every fp mul/add must round separately.
# TUNING=1 compiles the B200_EXP_* environment overrides in (experiments only; a release build reads none of them)
TUNEFLAG := $(if $(TUNING),-DB200_TUNING,) $(EXTRA_DEFS)
NVFLAGS := $(TUNEFLAG) $(ARCH) -O3 -std=c++17 -lineinfo -fmad=false -Xcompiler -fPIC,-O2,-fvisibility=hidden,-ffp-contract=off -Xptxas -v
CSRC := opencv_b200/csrc
OBJDIR := build/obj
LIB := opencv_b200/lib/libb200geom.so
SRCS := $(sort $(wildcard $(CSRC)/*.cu))
OBJS := $(patsubst $(CSRC)/%.cu,$(OBJDIR)/%.o,$(SRCS))
HDRS := $(wildcard $(CSRC)/*.cuh) include/b200_geom.h

all: $(LIB) oracle

$(OBJDIR)/%.o: $(CSRC)/%.cu $(HDRS)
	@mkdir -p $(OBJDIR)
	$(NVCC) $(NVFLAGS) -c $< -o $@ 2> $(OBJDIR)/$*.ptxas.log || (cat $(OBJDIR)/$*.ptxas.log; false)

$(LIB): $(OBJS)
	@mkdir -p opencv_b200/lib
	$(NVCC) $(ARCH) -shared -o $@ $(OBJS) -Xlinker --version-script=$(CSRC)/exports.map

oracle:
	$(MAKE) -C oracle -s

# the in-tree HAL route (hal/cuda_b200/CMakeLists.txt): configure + build with cmake and nvcc,
# then check that the result exports the C-ABI (run by __graft_entry__.build() so the recipe cannot rot)
cmake-check:
	cmake -S hal/cuda_b200 -B build/cmake_hal -DCMAKE_BUILD_TYPE=Release > build/cmake_hal.log 2>&1 || (cat build/cmake_hal.log; false)
	cmake --build build/cmake_hal -j >> build/cmake_hal.log 2>&1 || (tail -50 build/cmake_hal.log; false)
	python3 -c "import ctypes; l = ctypes.CDLL('build/cmake_hal/libb200geom.so'); [getattr(l, n) for n in ('b200_hal_warpAffine', 'b200_hal_resize', 'b200_hal_remap32f', 'b200_version')]; l.b200_version.restype = ctypes.c_char_p; print('cmake-check ok:', l.b200_version().decode())"

clean:
	rm -rf build $(LIB)

.PHONY: all oracle clean cmake-check
