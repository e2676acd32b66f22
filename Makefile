# Top-level build: libgi_b200.so (CUDA, sm_100a) + host tools.  `python -c "import __graft_entry__ as g; g.build()"`
# drives the same targets.
PKG    := global-illumination_b200
NVCC   ?= /usr/local/cuda/bin/nvcc
CXX    ?= g++
ARCH   := -gencode arch=compute_100a,code=sm_100a
# -fmad=false: parity with the reference's x86-64 build needs un-fused FP64 multiply/add
NVFLAGS := -O3 -std=c++17 $(ARCH) -lineinfo -fmad=false -Xcompiler -fPIC,-ffp-contract=off -Iinclude -I$(PKG)/host -I$(PKG)/csrc
CXXFLAGS := -O2 -std=c++17 -fPIC -Wall -Iinclude -I$(PKG)/host

HOST_SRCS := $(PKG)/host/scene_loader.cpp $(PKG)/host/scene_io.cpp
IMG_SRCS  := $(PKG)/host/image_io.cpp
CU_SRCS   := $(wildcard $(PKG)/csrc/*.cu)
CU_HDRS   := $(wildcard $(PKG)/csrc/*.cuh) $(wildcard $(PKG)/csrc/*.h) include/gi_b200.h $(PKG)/host/scene.h

.PHONY: all lib tools oracle host_sim clean
all: lib tools oracle host_sim

tools: $(PKG)/bin/gi_scene_tool $(PKG)/bin/photonmap
# the reference's program name; links against the C ABI only (rpath -> the in-tree library)
$(PKG)/bin/photonmap: $(PKG)/host/photonmap_main.cpp include/gi_b200.h $(PKG)/libgi_b200.so
	@mkdir -p $(PKG)/bin
	$(CXX) $(CXXFLAGS) $(PKG)/host/photonmap_main.cpp -o $@ -L$(PKG) -lgi_b200 -Wl,-rpath,'$$ORIGIN/..' -L/usr/local/cuda/lib64 -Wl,-rpath,/usr/local/cuda/lib64
$(PKG)/bin/gi_scene_tool: $(PKG)/host/scene_tool.cpp $(HOST_SRCS) $(PKG)/host/scene.h include/gi_b200.h
	@mkdir -p $(PKG)/bin
	$(CXX) $(CXXFLAGS) $< $(HOST_SRCS) -o $@

# one object per translation unit under build/ (git-ignored) so `make -j` compiles the kernels in parallel
LIB_SRCS := $(CU_SRCS) $(HOST_SRCS) $(IMG_SRCS) $(PKG)/host/capi_host.cpp
LIB_OBJS := $(patsubst %,build/%.o,$(notdir $(LIB_SRCS)))
vpath %.cu $(PKG)/csrc
vpath %.cpp $(PKG)/host
build/%.cu.o: %.cu $(CU_HDRS)
	@mkdir -p build
	$(NVCC) $(NVFLAGS) -c $< -o $@
build/%.cpp.o: %.cpp $(CU_HDRS)
	@mkdir -p build
	$(NVCC) $(NVFLAGS) -c $< -o $@

lib: $(PKG)/libgi_b200.so
$(PKG)/libgi_b200.so: $(LIB_OBJS)
	$(NVCC) $(ARCH) -shared $(LIB_OBJS) -o $@ -lcudart -ldl

host_sim: tests/host_sim/libisect_sim.so
tests/host_sim/libisect_sim.so: tests/host_sim/isect_sim.cpp $(CU_HDRS) $(HOST_SRCS)
	$(CXX) -O2 -std=c++17 -fPIC -shared -ffp-contract=off -w -Iinclude -I$(PKG)/host -I$(PKG)/csrc $< $(HOST_SRCS) -o $@ -lpthread

oracle:
	$(MAKE) -C oracle all

clean:
	rm -rf build $(PKG)/bin $(PKG)/libgi_b200.so
	$(MAKE) -C oracle clean
