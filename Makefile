# libthorcuda: hand-written CUDA for B200 (sm_100a) behind a C ABI.  `make` builds
# thor_b200/libthorcuda.so in-tree; `make oracle` builds the CPU checkers (test infrastructure).
NVCC      ?= nvcc
NVCCFLAGS ?= -O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -lineinfo -Xcompiler -fPIC -Xptxas -v
CSRC      := thor_b200/csrc
SRCS      := $(wildcard $(CSRC)/*.cu)
OBJS      := $(patsubst $(CSRC)/%.cu,build/%.o,$(SRCS))
LIB       := thor_b200/libthorcuda.so

.PHONY: all oracle clean
all: $(LIB)

build/%.o: $(CSRC)/%.cu $(CSRC)/internal.cuh $(CSRC)/lists.cuh include/thorcuda.h include/thorcuda_compat.h include/thorcuda_enc.h $(wildcard $(CSRC)/enc/*.h) | build
	$(NVCC) $(NVCCFLAGS) -c -o $@ $< 2> build/$*.ptxas.log || (cat build/$*.ptxas.log; false)

$(LIB): $(OBJS)
	$(NVCC) -shared -cudart static -o $@ $^

build:
	mkdir -p build

oracle:
	$(MAKE) -C oracle all

clean:
	rm -rf build $(LIB)
