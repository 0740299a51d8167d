# Builds libbrisk_b200.so (hand-written sm_100a CUDA + the C ABI) in-tree, and the test-only oracle.
NVCC      ?= /usr/local/cuda/bin/nvcc
ARCH      := -gencode arch=compute_100a,code=sm_100a
NVCCFLAGS := -O3 -std=c++17 -lineinfo $(ARCH) -Xcompiler -fPIC,-Wall,-Wno-unused-function -Xptxas -v
SRC       := brisk_b200/csrc/brisk_capi.cu
HDR       := $(wildcard brisk_b200/csrc/*.cuh) include/brisk_b200.h
LIB       := brisk_b200/libbrisk_b200.so

all: $(LIB) oracle

$(LIB): $(SRC) $(HDR)
	$(NVCC) $(NVCCFLAGS) -shared -o $@ $(SRC) -ldl 2> build_ptxas.log || (cat build_ptxas.log; exit 1)

oracle:
	$(MAKE) -s -C oracle liboracle.so

clean:
	rm -f $(LIB) build_ptxas.log
.PHONY: all oracle clean
