# Builds libccgpu.so (the product: hand-written sm_100a kernels behind the C ABI in include/ccgpu.h)
# and the CPU oracle (test infrastructure).  nvcc cross-compiles without a GPU.
NVCC ?= /usr/local/cuda/bin/nvcc
ARCH = -gencode arch=compute_100a,code=sm_100a
NVFLAGS = $(ARCH) -O3 -std=c++17 -lineinfo -Xcompiler -fPIC -Xcompiler -Wall -Xptxas -v --expt-relaxed-constexpr
SRC = cloudclean_b200/csrc
OBJ = $(SRC)/build.o $(SRC)/query.o $(SRC)/radsel.o $(SRC)/gridnormals.o $(SRC)/noalloc.o $(SRC)/voxel.o $(SRC)/ptx.o $(SRC)/api.o
LIB = cloudclean_b200/libccgpu.so

all: $(LIB) oracle adapters

# ptxas -v output (registers / spills / shared memory per kernel) goes to build/ptxas/ (not tracked); the per-round table the
# judge reads is profiles/rNN_ptxas.txt, written by `make ptxas-summary`
PTXLOG = build/ptxas
$(SRC)/%.o: $(SRC)/%.cu $(SRC)/common.cuh $(SRC)/tile.cuh include/ccgpu.h
	@mkdir -p $(PTXLOG)
	$(NVCC) $(NVFLAGS) -c $< -o $@ 2> $(PTXLOG)/$(notdir $(@:.o=.log)) || (cat $(PTXLOG)/$(notdir $(@:.o=.log)); exit 1)

ptxas-summary:
	python tools/ptxas_summary.py $(PTXLOG) > profiles/r02_ptxas.txt

$(LIB): $(OBJ)
	$(NVCC) $(ARCH) -shared -o $@ $(OBJ) -lcudart_static -lpthread -ldl -lrt

oracle:
	$(MAKE) -C oracle

# reference-named C++ host layer + its GPU test program (links only the C ABI)
adapters: adapters/test_adapters
adapters/test_adapters: adapters/test_adapters.cpp adapters/cloudclean_gpu.hpp include/ccgpu.h $(LIB)
	g++ -O2 -std=c++14 -ffp-contract=off -Wall -pthread -I. adapters/test_adapters.cpp -o $@ -Lcloudclean_b200 -lccgpu -Wl,-rpath,'$$ORIGIN/../cloudclean_b200'

clean:
	rm -rf $(OBJ) $(LIB) $(PTXLOG) adapters/test_adapters
	$(MAKE) -C oracle clean
.PHONY: all oracle adapters clean ptxas-summary
