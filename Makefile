# Builds libpeasoup_b200.so (hand-written sm_100a kernels + C ABI + C++ pipeline driver).
# Test infrastructure (CPU oracle, reference harness) is built by oracle/Makefile.
NVCC ?= nvcc
ARCH := -gencode arch=compute_100a,code=sm_100a
NVFLAGS := $(ARCH) $(EXTRA) -O3 -std=c++17 -lineinfo -Xcompiler -fPIC,-Wall,-Wno-unused-function -Iinclude
CSRC := peasoup_b200/csrc
LIB := peasoup_b200/lib/libpeasoup_b200.so
OBJS := build/stages.o build/fft.o build/bluestein.o build/search.o build/dedisp.o build/fold.o build/coincidence.o build/pipeline.o

all: $(LIB)

build/%.o: $(CSRC)/%.cu $(wildcard $(CSRC)/*.cuh) $(wildcard include/*.h) $(wildcard include/peasoup_b200/*.hpp)
	@mkdir -p build
	$(NVCC) $(NVFLAGS) -Xptxas -v -c $< -o $@ 2> build/$*.ptxas.log || (cat build/$*.ptxas.log; false)

build/pipeline.o: $(CSRC)/pipeline.cpp $(wildcard include/*.h) $(wildcard include/peasoup_b200/*.hpp)
	@mkdir -p build
	$(NVCC) $(NVFLAGS) -x cu -c $< -o $@

$(LIB): $(OBJS)
	@mkdir -p peasoup_b200/lib
	$(NVCC) $(ARCH) -shared -o $@ $(OBJS) -lcudart

oracle:
	$(MAKE) -C oracle all

clean:
	rm -rf build peasoup_b200/lib

.PHONY: all oracle clean
