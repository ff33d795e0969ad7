# Top-level build: CUDA engine (libfeelgpu.so), host-side mirror (libfeelhost.so), oracle (test infrastructure).
NVCC ?= /usr/local/cuda/bin/nvcc
CXX ?= g++
ARCH := -gencode arch=compute_100a,code=sm_100a
EPLAN_DEFS ?=
NVCCFLAGS := $(ARCH) -O3 -lineinfo -std=c++17 -Xcompiler -fPIC -Xptxas -v $(EPLAN_DEFS)
CSRC := feelpp_b200/csrc
LIB := feelpp_b200/lib

all: $(LIB)/libfeelgpu.so $(LIB)/libfeelhost.so oracle

$(LIB)/%.o: $(CSRC)/%.cu $(CSRC)/feelgpu_internal.h $(CSRC)/rowplan.h $(CSRC)/eplan.h $(CSRC)/rplan.h $(CSRC)/device_common.cuh include/feelgpu.h
	@mkdir -p $(LIB)
	$(NVCC) $(NVCCFLAGS) -c $< -o $@ 2> $(LIB)/$*.ptxas.log || (cat $(LIB)/$*.ptxas.log; false)

$(LIB)/rowplan_build.o: $(CSRC)/rowplan_build.cpp $(CSRC)/rowplan.h
	@mkdir -p $(LIB)
	$(CXX) -O3 -march=x86-64-v3 -std=c++17 -fPIC -Wall -pthread -c $< -o $@

$(LIB)/rplan_build.o: $(CSRC)/rplan_build.cpp $(CSRC)/rplan.h
	@mkdir -p $(LIB)
	$(CXX) -O3 -march=x86-64-v3 -std=c++17 -fPIC -Wall -pthread -c $< -o $@

$(LIB)/eplan_build.o: $(CSRC)/eplan_build.cpp $(CSRC)/eplan.h
	@mkdir -p $(LIB)
	$(CXX) -O3 -march=x86-64-v3 -std=c++17 -fPIC -Wall -pthread $(EPLAN_DEFS) -c $< -o $@

# NCCL (exchange.cu) is bound with dlopen at feelgpu_comm_init, not linked: see the note at the top of exchange.cu
$(LIB)/libfeelgpu.so: $(LIB)/feelgpu_api.o $(LIB)/kernels_scatter.o $(LIB)/kernels_rowplan.o $(LIB)/rowplan_build.o $(LIB)/kernels_eplan.o $(LIB)/eplan_build.o $(LIB)/kernels_on.o $(LIB)/kernels_graph.o $(LIB)/exchange.o $(LIB)/kernels_rplan.o $(LIB)/rplan_build.o
	$(NVCC) $(ARCH) -shared -o $@ $^ -lcudart -ldl

# METIS 5.1.0 as vendored by the reference (feelpp/contrib/metis: the partitioner Feel++ itself uses,
# feelpartition/partitionermetis_impl.hpp).  Its C sources are compiled where they lie under /root/reference into a static
# archive (a built artefact like the .so files: git-ignored, it travels to the GPU box; no source is copied).  Where the
# reference tree is absent (the GPU box) the archive built here is used as it is; without either, the host mirror is
# built without METIS and the harness partitions by coordinate bisection.
REF_METIS ?= /root/reference/feelpp/contrib/metis
METIS_A := $(LIB)/libfeelmetis.a
ifneq ($(wildcard $(REF_METIS)/libmetis/pmetis.c),)
METIS_SRC := $(wildcard $(REF_METIS)/GKlib/*.c) $(wildcard $(REF_METIS)/libmetis/*.c)
$(METIS_A): $(METIS_SRC)
	@mkdir -p $(LIB)/metis_obj
	@for f in $(METIS_SRC); do echo "gcc -O2 -fPIC -DLINUX -D_FILE_OFFSET_BITS=64 -std=c99 -fno-strict-aliasing -DNDEBUG -DNDEBUG2 -DHAVE_EXECINFO_H -DHAVE_GETLINE -w -I$(REF_METIS)/GKlib -I$(REF_METIS)/include -I$(REF_METIS)/libmetis -c $$f -o $(LIB)/metis_obj/`basename $$(dirname $$f)`_`basename $$f .c`.o"; done | xargs -P 8 -I CMD sh -c CMD
	rm -f $@ && ar rcs $@ $(LIB)/metis_obj/*.o
HAVE_METIS := 1
else ifneq ($(wildcard $(METIS_A)),)
HAVE_METIS := 1
endif
ifeq ($(HAVE_METIS),1)
METIS_LINK := $(METIS_A)
METIS_DEF :=
else
METIS_LINK :=
METIS_DEF := -DFEELHOST_NO_METIS
endif

HOSTSRC := $(wildcard feelpp_b200/host/*.cpp)
HOSTHDR := $(wildcard feelpp_b200/host/*.hpp) include/feelgpu.h include/feelgpu_quadrature_tables.h
$(LIB)/libfeelhost.so: $(HOSTSRC) $(HOSTHDR) $(METIS_LINK)
	@mkdir -p $(LIB)
	$(CXX) -O3 -march=x86-64-v3 -std=c++17 -fPIC -Wall -pthread -shared -Iinclude $(METIS_DEF) -o $@ $(HOSTSRC) $(METIS_LINK) -lm

oracle:
	$(MAKE) -C oracle

clean:
	rm -rf $(LIB) oracle/_build

.PHONY: all oracle clean
