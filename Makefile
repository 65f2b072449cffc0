# Builds the product: libgoofit_b200.so (hand-written sm_100a kernels behind the C ABI in include/goofit_b200.h)
NVCC     ?= /usr/local/cuda/bin/nvcc
ARCH     := -gencode arch=compute_100a,code=sm_100a
NVFLAGS  := -O3 -std=c++17 $(ARCH) -lineinfo -Xcompiler -fPIC -Xptxas -v --expt-relaxed-constexpr
CSRC     := goofit_b200/csrc
LIBDIR   := goofit_b200/lib
CU       := $(wildcard $(CSRC)/*.cu)
OBJ      := $(patsubst $(CSRC)/%.cu,build/%.o,$(CU))

all: $(LIBDIR)/libgoofit_b200.so

build/%.o: $(CSRC)/%.cu $(wildcard $(CSRC)/*.cuh) include/goofit_b200.h
	@mkdir -p build
	$(NVCC) $(NVFLAGS) -c $< -o $@ 2> build/$*.ptxas.log || (cat build/$*.ptxas.log; false)

$(LIBDIR)/libgoofit_b200.so: $(OBJ)
	@mkdir -p $(LIBDIR)
	$(NVCC) -shared $(ARCH) -o $@ $(OBJ) -lcudart

clean:
	rm -rf build $(LIBDIR)/*.so
.PHONY: all clean
