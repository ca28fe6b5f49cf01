# Builds libvio360_cuda.so (sm_100a only) in-tree so that it travels with gpurun snapshots.
NVCC ?= /usr/local/cuda/bin/nvcc
PKG := 360_visual_inertial_odometry_b200
CSRC := $(PKG)/csrc
ARCH := -gencode arch=compute_100a,code=sm_100a
NVFLAGS := $(ARCH) -O3 -lineinfo -std=c++17 -Xcompiler -fPIC -Xcompiler -Wall -Xptxas -v $(EXTRA_NVFLAGS)
LIB := $(PKG)/libvio360_cuda.so
OBJS := $(CSRC)/ctx.o $(CSRC)/ba.o $(CSRC)/fe.o

all: $(LIB) oracle host hosttest

$(CSRC)/ctx.o: $(CSRC)/ctx.cu $(CSRC)/common.cuh include/vio360.h
	$(NVCC) $(NVFLAGS) -c $< -o $@
$(CSRC)/ba.o: $(CSRC)/ba.cu $(CSRC)/ba_peer.cuh $(CSRC)/ba_math.cuh $(CSRC)/ba_schur.cuh $(CSRC)/ba_schur_kernels.cuh $(CSRC)/ba_solve.cuh $(CSRC)/ba_lm.cuh $(CSRC)/ba_triangulate.cuh $(CSRC)/common.cuh include/vio360.h
	$(NVCC) $(NVFLAGS) -c $< -o $@
# --fmad=false: float expressions must round like OpenCV's non-FMA baseline code; the
# places where OpenCV's dispatched code does fuse are explicit fmaf() calls.
$(CSRC)/fe.o: $(CSRC)/fe.cu $(CSRC)/fe_seq.cuh $(CSRC)/fe_seq_host.inc $(CSRC)/fe_grid.cuh $(CSRC)/fe_pyr.cuh $(CSRC)/fe_lk.cuh $(CSRC)/fe_detect.cuh $(CSRC)/fe_ransac.cuh $(CSRC)/common.cuh include/vio360.h
	$(NVCC) $(NVFLAGS) --fmad=false -c $< -o $@

$(LIB): $(OBJS)
	$(NVCC) $(ARCH) -shared -o $@ $(OBJS) -lcudart

oracle:
	$(MAKE) -C oracle

host: $(LIB)
	@if [ -f $(PKG)/host/Makefile ]; then $(MAKE) -C $(PKG)/host; fi

clean:
	rm -f $(OBJS) $(LIB)
	$(MAKE) -C oracle clean

hosttest: host oracle
	$(MAKE) -C tests/host

.PHONY: all oracle host hosttest clean
