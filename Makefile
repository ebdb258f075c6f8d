# Top-level build.  Everything is built IN-TREE so that the shared objects
# travel to the GPU box with the repository snapshot.
#
#   simple_raytracer_b200/lib/librt_b200.so   CUDA render core (C ABI, include/rt.h), sm_100a
#   simple_raytracer_b200/lib/librt_host.so   CPU front-end (parser, P3 reader/writer; include/rt_host.h)
#   bin/raytracer                             `raytracer scene.txt` CLI
#   oracle/liboracle.so, oracle/_ref/*        test oracle + the reference compiled from /root/reference
PKG   := simple_raytracer_b200
CSRC  := $(PKG)/csrc
LIB   := $(PKG)/lib
NVCC  ?= /usr/local/cuda/bin/nvcc
CXX   ?= g++

ARCH      := -gencode arch=compute_100a,code=sm_100a
# -fmad is left at its default: every result-defining operation in the kernels
# goes through explicit __f*_rn / __d*_rn intrinsics (never contracted); only
# conservative culling arithmetic may fuse.  No -use_fast_math.
NVCCFLAGS := -std=c++17 -O3 $(ARCH) -lineinfo -Xcompiler -fPIC,-ffp-contract=off -Iinclude -I$(CSRC) \
             --expt-relaxed-constexpr -cudart static
HOSTFLAGS := -std=c++17 -O2 -ffp-contract=off -fPIC -pthread -Wall -Wextra -Iinclude -I$(CSRC)/host

CU_SRCS   := $(wildcard $(CSRC)/*.cu)
CU_HDRS   := $(wildcard $(CSRC)/*.cuh) $(wildcard $(CSRC)/*.h) include/rt.h
HOST_SRCS := $(CSRC)/host/scene_parser.cpp $(CSRC)/host/ppm_io.cpp

all: $(LIB)/librt_host.so $(LIB)/librt_b200.so bin/raytracer oracle

$(LIB)/librt_host.so: $(HOST_SRCS) $(CSRC)/host/host_common.h include/rt_host.h include/rt.h
	@mkdir -p $(LIB)
	$(CXX) $(HOSTFLAGS) -shared -o $@ $(HOST_SRCS)

# one object per .cu so that `make -j` compiles them side by side (render.cu alone is most of the time)
CU_OBJS   := $(patsubst $(CSRC)/%.cu,build/%.o,$(CU_SRCS))

build/%.o: $(CSRC)/%.cu $(CU_HDRS)
	@mkdir -p build $(LIB)
	$(NVCC) $(NVCCFLAGS) $(EXTRA) -Xptxas -v -c -o $@ $< 2> build/$*.ptxas.log || (cat build/$*.ptxas.log; exit 1)

$(LIB)/librt_b200.so: $(CU_OBJS)
	@mkdir -p $(LIB)
	$(NVCC) $(ARCH) -cudart static -shared -o $@ $(CU_OBJS) -ldl
	@cat build/*.ptxas.log > $(LIB)/ptxas.log

bin/raytracer: $(CSRC)/host/raytracer_main.cpp $(LIB)/librt_host.so $(LIB)/librt_b200.so
	@mkdir -p bin
	$(CXX) -std=c++17 -O2 -Iinclude -o $@ $< -L$(LIB) -lrt_b200 -lrt_host -Wl,-rpath,'$$ORIGIN/../$(LIB)'

oracle: $(LIB)/librt_b200.so
	$(MAKE) -C oracle all

clean:
	rm -rf $(LIB) bin build
	$(MAKE) -C oracle clean

.PHONY: all oracle clean
