# Builds everything in-tree (the .so files travel to the GPU box with the snapshot).
#   librrt.so       CUDA kernels + C ABI (include/rrt.h), sm_100a only
#   librrt_host.so  host mirror of the reference's Rust surface (include/rrt_host.h)
#   bin/rrt         CLI mirroring src/main.rs
#   oracle          CPU restatement (test infrastructure) -> oracle/liboracle.so
NVCC      ?= nvcc
CXX       ?= g++
PKG       := rustraytrace_b200
NVFLAGS   := -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 --use_fast_math \
             -Xcompiler -fPIC -Xcompiler -Wall
CXXFLAGS  := -O2 -std=c++17 -fPIC -Wall

CSRC := $(PKG)/csrc
HOST := $(PKG)/host

all: $(PKG)/librrt.so $(PKG)/librrt_host.so $(PKG)/bin/rrt oracle

$(CSRC)/kernels.o: $(CSRC)/kernels.cu $(CSRC)/*.cuh $(CSRC)/*.h include/rrt.h
	$(NVCC) $(NVFLAGS) -c $< -o $@
# render_kernel<INTEGRATOR, FEAT> instances, one translation unit per book so they build in parallel
$(CSRC)/kernels_book%.o: $(CSRC)/kernels_book%.cu $(CSRC)/*.cuh $(CSRC)/*.h include/rrt.h
	$(NVCC) $(NVFLAGS) -c $< -o $@
$(CSRC)/probe64.o: $(CSRC)/probe64.cu $(CSRC)/*.cuh $(CSRC)/*.h include/rrt.h
	$(NVCC) -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -fmad=false -Xcompiler -fPIC -c $< -o $@
$(CSRC)/api.o: $(CSRC)/api.cu $(CSRC)/*.cuh $(CSRC)/*.h $(CSRC)/flatten.hpp include/rrt.h
	$(NVCC) $(NVFLAGS) -c $< -o $@
# several GPUs behind the C ABI: NCCL reduce / gather of the frame (links libnccl)
$(CSRC)/multi.o: $(CSRC)/multi.cu $(CSRC)/*.h $(CSRC)/flatten.hpp include/rrt.h
	$(NVCC) $(NVFLAGS) -c $< -o $@
# rrt_probe_static: the probe with its scratch in a static device array (no cudaMalloc)
$(CSRC)/noalloc.o: $(CSRC)/noalloc.cu $(CSRC)/*.h $(CSRC)/flatten.hpp include/rrt.h
	$(NVCC) $(NVFLAGS) -c $< -o $@
$(CSRC)/flatten.o: $(CSRC)/flatten.cpp $(CSRC)/flatten.hpp $(CSRC)/device_types.h include/rrt.h
	$(NVCC) $(NVFLAGS) -x cu -c $< -o $@

KOBJ := $(CSRC)/kernels_book1.o $(CSRC)/kernels_book2.o $(CSRC)/kernels_book3.o
$(PKG)/librrt.so: $(CSRC)/kernels.o $(KOBJ) $(CSRC)/probe64.o $(CSRC)/api.o $(CSRC)/multi.o $(CSRC)/noalloc.o $(CSRC)/flatten.o
	$(NVCC) -gencode arch=compute_100a,code=sm_100a -shared -o $@ $^ -cudart shared -lnccl

HOST_SRC := $(HOST)/books.cpp $(HOST)/rtw_image.cpp $(HOST)/jpeg_decode.cpp $(HOST)/render_io.cpp
$(PKG)/librrt_host.so: $(HOST_SRC) $(HOST)/*.hpp include/*.h
	$(CXX) $(CXXFLAGS) -shared -o $@ $(HOST_SRC)

$(PKG)/bin/rrt: $(HOST)/main.cpp $(PKG)/librrt.so $(PKG)/librrt_host.so
	mkdir -p $(PKG)/bin
	$(CXX) $(CXXFLAGS) -o $@ $(HOST)/main.cpp -L$(PKG) -lrrt -lrrt_host -Wl,-rpath,'$$ORIGIN/..'

oracle:
	$(MAKE) -C oracle

# CPU model of the warp-level traversal cost (scripts/warp_sim.cpp; not part of `all`, not shipped):
#   make warp_sim [SIMFLAGS="-DRRT_FRONT_MAX=8"] && build/warp_sim 1 -1 1200
warp_sim: $(PKG)/librrt_host.so
	mkdir -p build
	$(CXX) -O2 -std=c++17 $(SIMFLAGS) -I$(CSRC) scripts/warp_sim.cpp $(CSRC)/flatten.cpp -L$(PKG) -lrrt_host -Wl,-rpath,$(CURDIR)/$(PKG) -o build/warp_sim

clean:
	rm -f $(CSRC)/*.o $(PKG)/*.so $(PKG)/bin/rrt
	$(MAKE) -C oracle clean

.PHONY: all oracle clean warp_sim
