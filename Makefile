# Builds libb200geom.so (sm_100a only) and the test oracle.
NVCC ?= nvcc
ARCH := -gencode arch=compute_100a,code=sm_100a
# -fmad=false: the reference's default build has no FMA; every fp mul/add must round separately.
NVFLAGS := $(ARCH) -O3 -std=c++17 -lineinfo -fmad=false -Xcompiler -fPIC,-O2,-fvisibility=hidden,-ffp-contract=off -Xptxas -v
CSRC := opencv_b200/csrc
OBJDIR := build/obj
LIB := opencv_b200/lib/libb200geom.so
SRCS := $(CSRC)/interp_tables.cu $(CSRC)/warp_kernels.cu $(CSRC)/resize_kernels.cu $(CSRC)/fast_kernels.cu $(CSRC)/capi.cu $(CSRC)/noalloc_pipeline.cu
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

clean:
	rm -rf build $(LIB)

.PHONY: all oracle clean
