# Builds everything in-tree (the .so files travel to the GPU box with the snapshot).
#   make            host library, oracle, CUDA library, harness CLI
#   make gpu        only libvgaligner_gpu.so (nvcc cross-compiles sm_100a without a GPU)
NVCC      ?= nvcc
CXX       ?= g++
ARCH      := -gencode arch=compute_100a,code=sm_100a
NVFLAGS   := $(ARCH) -lineinfo -O3 -std=c++17 -Xcompiler -fPIC -Xcompiler -Wall
CXXFLAGS  := -O2 -std=c++17 -fPIC -Wall -Wextra
PKG       := rs-vgaligner_b200
CSRC      := $(PKG)/csrc
HOST      := $(PKG)/host
BUILD     := build
GPU_OBJS  := $(BUILD)/vga_multi.o $(BUILD)/vga_index.o $(BUILD)/vga_api.o $(BUILD)/k_seed.o $(BUILD)/k_chain.o $(BUILD)/k_subgraph.o $(BUILD)/k_poa.o $(BUILD)/k_poa_short.o $(BUILD)/k_poa_pair.o $(BUILD)/k_poa_long.o $(BUILD)/k_poa_wide.o $(BUILD)/k_index_build.o $(BUILD)/vga_noalloc.o

all: host oracle gpu cli

$(BUILD):
	mkdir -p $(BUILD)

# ---- CUDA library (the product) ----
gpu: $(PKG)/libvgaligner_gpu.so
$(BUILD)/k_chain.o: $(CSRC)/k_chain.cu $(CSRC)/common.cuh $(CSRC)/internal.h include/vga.h | $(BUILD)
	$(NVCC) $(NVFLAGS) -fmad=false -c $< -o $@
$(BUILD)/%.o: $(CSRC)/%.cu $(CSRC)/common.cuh $(CSRC)/internal.h include/vga.h | $(BUILD)
	$(NVCC) $(NVFLAGS) -c $< -o $@
$(PKG)/libvgaligner_gpu.so: $(GPU_OBJS)
	$(NVCC) $(ARCH) -shared -o $@ $(GPU_OBJS) -ldl

# ---- host library (index build, reads, GAF writer) ----
host: $(PKG)/libvga_host.so
HOST_SRC  := $(HOST)/index_build.cpp $(HOST)/host_api.cpp $(HOST)/gaf_writer.cpp $(HOST)/key_derive.cpp $(HOST)/idx_file.cpp
$(PKG)/libvga_host.so: $(HOST_SRC) $(HOST)/flat_index.h $(HOST)/vga_host.h include/vga.h
	$(CXX) $(CXXFLAGS) -O3 -pthread -shared -o $@ $(HOST_SRC)

# ---- harness CLI mirroring `vgaligner index` / `vgaligner map` ----
cli: $(PKG)/vga_map
$(PKG)/vga_map: $(HOST)/vga_map_main.cpp $(PKG)/libvga_host.so $(PKG)/libvgaligner_gpu.so
	$(CXX) $(CXXFLAGS) -o $@ $(HOST)/vga_map_main.cpp -L$(PKG) -lvga_host -lvgaligner_gpu -Wl,-rpath,'$$ORIGIN'

# ---- oracle (test infrastructure only) ----
oracle: oracle/_build/liboracle.so
oracle/_build/liboracle.so: oracle/oracle.cpp oracle/oracle.h
	mkdir -p oracle/_build
	$(CXX) -O2 -std=c++17 -ffp-contract=off -pthread -fPIC -Wall -Wextra -shared -o $@ oracle/oracle.cpp

clean:
	rm -rf $(BUILD) $(PKG)/*.so $(PKG)/vga_map oracle/_build

.PHONY: all gpu host cli oracle clean
