# Builds libccgpu.so (the product: hand-written sm_100a kernels behind the C ABI in include/ccgpu.h)
# and the CPU oracle (test infrastructure).  nvcc cross-compiles without a GPU.
NVCC ?= /usr/local/cuda/bin/nvcc
ARCH = -gencode arch=compute_100a,code=sm_100a
NVFLAGS = $(ARCH) -O3 -std=c++17 -lineinfo -Xcompiler -fPIC -Xcompiler -Wall -Xptxas -v --expt-relaxed-constexpr
SRC = cloudclean_b200/csrc
OBJ = $(SRC)/build.o $(SRC)/query.o $(SRC)/gridnormals.o $(SRC)/api.o
LIB = cloudclean_b200/libccgpu.so

all: $(LIB) oracle

$(SRC)/%.o: $(SRC)/%.cu $(SRC)/common.cuh include/ccgpu.h
	$(NVCC) $(NVFLAGS) -c $< -o $@ 2> $(@:.o=.ptxas.log) || (cat $(@:.o=.ptxas.log); exit 1)

$(LIB): $(OBJ)
	$(NVCC) $(ARCH) -shared -o $@ $(OBJ) -lcudart_static -lpthread -ldl -lrt

oracle:
	$(MAKE) -C oracle

clean:
	rm -f $(OBJ) $(LIB) $(SRC)/*.ptxas.log
	$(MAKE) -C oracle clean
.PHONY: all oracle clean
