# Top-level build: CUDA engine (libfeelgpu.so), host-side mirror (libfeelhost.so), oracle (test infrastructure).
NVCC ?= /usr/local/cuda/bin/nvcc
CXX ?= g++
ARCH := -gencode arch=compute_100a,code=sm_100a
EPLAN_DEFS ?=
NVCCFLAGS := $(ARCH) -O3 -lineinfo -std=c++17 -Xcompiler -fPIC -Xptxas -v $(EPLAN_DEFS)
CSRC := feelpp_b200/csrc
LIB := feelpp_b200/lib

all: $(LIB)/libfeelgpu.so $(LIB)/libfeelhost.so oracle

$(LIB)/%.o: $(CSRC)/%.cu $(CSRC)/feelgpu_internal.h $(CSRC)/rowplan.h $(CSRC)/eplan.h include/feelgpu.h
	@mkdir -p $(LIB)
	$(NVCC) $(NVCCFLAGS) -c $< -o $@ 2> $(LIB)/$*.ptxas.log || (cat $(LIB)/$*.ptxas.log; false)

$(LIB)/rowplan_build.o: $(CSRC)/rowplan_build.cpp $(CSRC)/rowplan.h
	@mkdir -p $(LIB)
	$(CXX) -O3 -march=x86-64-v3 -std=c++17 -fPIC -Wall -pthread -c $< -o $@

$(LIB)/eplan_build.o: $(CSRC)/eplan_build.cpp $(CSRC)/eplan.h
	@mkdir -p $(LIB)
	$(CXX) -O3 -march=x86-64-v3 -std=c++17 -fPIC -Wall -pthread $(EPLAN_DEFS) -c $< -o $@

$(LIB)/libfeelgpu.so: $(LIB)/feelgpu_api.o $(LIB)/kernels_scatter.o $(LIB)/kernels_rowplan.o $(LIB)/rowplan_build.o $(LIB)/kernels_eplan.o $(LIB)/eplan_build.o $(LIB)/kernels_on.o $(LIB)/kernels_graph.o
	$(NVCC) $(ARCH) -shared -o $@ $^ -lcudart

HOSTSRC := $(wildcard feelpp_b200/host/*.cpp)
HOSTHDR := $(wildcard feelpp_b200/host/*.hpp) include/feelgpu.h include/feelgpu_quadrature_tables.h
$(LIB)/libfeelhost.so: $(HOSTSRC) $(HOSTHDR)
	@mkdir -p $(LIB)
	$(CXX) -O3 -march=x86-64-v3 -std=c++17 -fPIC -Wall -pthread -shared -Iinclude -o $@ $(HOSTSRC)

oracle:
	$(MAKE) -C oracle

clean:
	rm -rf $(LIB) oracle/_build

.PHONY: all oracle clean
