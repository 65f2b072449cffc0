# Builds the product:
#   goofit_b200/lib/libgoofit_b200.so   hand-written sm_100a kernels behind the C ABI (include/goofit_b200.h)
#   goofit_b200/lib/libMinuit2.so       the reference's bundled third-party Minuit2, compiled unmodified from
#                                       $(REF)/extern/Minuit2 (host-side minimiser; needs the reference checkout)
#   goofit_b200/lib/libgoofit_host.so   host C++ API (GooFit class names) over the C ABI
#   goofit_b200/_goofit*.so             pybind11 module
NVCC     ?= /usr/local/cuda/bin/nvcc
CXX      ?= g++
PYTHON   ?= python
REF      ?= /root/reference
ARCH     := -gencode arch=compute_100a,code=sm_100a
NVFLAGS  := -O3 -std=c++17 $(ARCH) -lineinfo -Xcompiler -fPIC -Xptxas -v --expt-relaxed-constexpr
CSRC     := goofit_b200/csrc
HOST     := goofit_b200/host
LIBDIR   := goofit_b200/lib
CU       := $(wildcard $(CSRC)/*.cu)
OBJ      := $(patsubst $(CSRC)/%.cu,build/%.o,$(CU))
M2INC    := $(REF)/extern/Minuit2/inc
M2SRC    := $(filter-out %/TMinuit2TraceObject.cxx,$(wildcard $(REF)/extern/Minuit2/src/*.cxx) $(wildcard $(REF)/extern/Minuit2/src/math/*.cxx))
M2OBJ    := $(patsubst $(REF)/extern/Minuit2/src/%.cxx,build/minuit2/%.o,$(M2SRC))
HSRC     := $(wildcard $(HOST)/src/*.cpp)
HOBJ     := $(patsubst $(HOST)/src/%.cpp,build/host/%.o,$(HSRC))
HFLAGS   := -O2 -std=c++17 -fPIC -Wall -Wno-unused-function -Iinclude -I$(HOST)/include -I$(M2INC) -DMATHCORE_STANDALONE
# This image's g++ wrapper points -lstdc++ at a dangling symlink and silently falls back to the static libstdc++.a,
# which must never be embedded in several shared objects of one process: link the system's shared libstdc++ explicitly.
SYSGCC   := $(firstword $(wildcard /usr/lib/gcc/x86_64-linux-gnu/*/libstdc++.so))
LDSTD    := $(if $(SYSGCC),-L$(dir $(SYSGCC)),)
PYEXT    := $(shell $(PYTHON) -c "import sysconfig; print(sysconfig.get_config_var('EXT_SUFFIX'))")
PYINC    := $(shell $(PYTHON) -c "import sysconfig, pybind11; print('-I' + sysconfig.get_paths()['include'] + ' -I' + pybind11.get_include())")

.SECONDEXPANSION:
all: cuda host python examples
cuda: $(LIBDIR)/libgoofit_b200.so
host: $(LIBDIR)/libgoofit_host.so
python: goofit_b200/_goofit$(PYEXT)

build/%.o: $(CSRC)/%.cu $(wildcard $(CSRC)/*.cuh) include/goofit_b200.h
	@mkdir -p build
	$(NVCC) $(NVFLAGS) -c $< -o $@ 2> build/$*.ptxas.log || (cat build/$*.ptxas.log; false)

$(LIBDIR)/libgoofit_b200.so: $(OBJ)
	@mkdir -p $(LIBDIR)
	$(NVCC) -shared $(ARCH) -o $@ $(OBJ) -lcudart -ldl

build/minuit2/%.o: $(REF)/extern/Minuit2/src/%.cxx
	@mkdir -p $(dir $@)
	$(CXX) -std=c++17 -O2 -fPIC -w -DMATHCORE_STANDALONE -DROOT_Math_VecTypes -DMATH_NO_PLUGIN_MANAGER -I$(M2INC) -c $< -o $@

$(LIBDIR)/libMinuit2.so: $(M2OBJ)
	@mkdir -p $(LIBDIR)
	@if [ -n "$(M2OBJ)" ]; then $(CXX) -shared $(LDSTD) -o $@ $(M2OBJ); else test -f $@ || (echo "Minuit2 sources not found under $(REF) and no prebuilt $@"; false); fi

build/host/%.o: $(HOST)/src/%.cpp $(wildcard $(HOST)/include/goofit/detail/*.h) include/goofit_b200.h
	@mkdir -p build/host
	$(CXX) $(HFLAGS) -c $< -o $@

$(LIBDIR)/libgoofit_host.so: $(HOBJ) $(LIBDIR)/libgoofit_b200.so $(LIBDIR)/libMinuit2.so
	$(CXX) -shared $(LDSTD) -o $@ $(HOBJ) -L$(LIBDIR) -lgoofit_b200 -lMinuit2 -Wl,-rpath,'$$ORIGIN'

build/py/%.o: goofit_b200/python/%.cpp $(wildcard $(HOST)/include/goofit/detail/*.h)
	@mkdir -p build/py
	$(CXX) $(HFLAGS) $(PYINC) -fvisibility=hidden -c $< -o $@

PYSRC := $(wildcard goofit_b200/python/*.cpp)
PYOBJ := $(patsubst goofit_b200/python/%.cpp,build/py/%.o,$(PYSRC))
goofit_b200/_goofit$(PYEXT): $(PYOBJ) $(LIBDIR)/libgoofit_host.so
	$(CXX) -shared $(LDSTD) -o $@ $(PYOBJ) -L$(LIBDIR) -lgoofit_host -lgoofit_b200 -lMinuit2 -Wl,-rpath,'$$ORIGIN/lib'

# The reference's OWN example sources, compiled unmodified where they lie against this repo's headers and libraries (the ones that
# need no ROOT): what a user does who switches the include and library paths. Only where the reference checkout exists.
REFEX    := $(if $(wildcard $(REF)/examples/exponential/exponential.cpp),examples/bin/ref_exponential examples/bin/ref_product,)
# The reference's OWN Catch2 test sources, compiled unmodified where they lie and linked against this repo's libraries: the tests its
# tests/{convert,simple,PDFs}/CMakeLists.txt enable without ROOT, plus the ROOT-gated ones that never touch ROOT. Catch2 itself is an
# empty submodule upstream; tests/ref_catch provides the part of its interface these sources use. {fmt} comes from the reference's extern/.
RT_CONVERT := BifurGaussTest CompositeTest ConvolutionTest EventWeightedAddTest GaussianTest JohnsonSUTest KinLimitBWTest LandauTest \
              NovosibirskTest PolynomialTest ProdTest ScaledGaussianTest SmoothHistogramTest StepTest VoigtianTest BWTest CorrGaussianTest
RT_SIMPLE  := NormalizeTest AddNormTest SimpleTest BinningTest BlindTest MonteCarloTest
RT_PDFS    := GenArgusTest GenGaussianTest
# Minuit1Test fits through goofit/fitting/FitManagerMinuit1.h: GOOFIT_ROOT_FOUND=1 with tests/ref_root, whose TMinuit stand-in is
# carried by Minuit2. The five convert tests that call test_fitter_minuit1 (Argus, BinTransform, CrystalBall, ExpGauss, Exp) compile
# the same way but are not built: they are degenerate fits (two free parameters for one exponential slope, a PDF without
# parameters) whose expectations are the point where the Fortran-descended TMinuit happens to stop, which no stand-in reproduces.
RT_ROOT_CONVERT :=
RT_ROOT_SIMPLE  := Minuit1Test
RT_ROOT_OBJ := $(RT_ROOT_CONVERT:%=build/reftests/convert/%.o) $(RT_ROOT_SIMPLE:%=build/reftests/simple/%.o)
$(RT_ROOT_OBJ): RT_EXTRA := -DGOOFIT_ROOT_FOUND=1 -Itests/ref_root
$(RT_ROOT_OBJ): $(wildcard tests/ref_root/*.h) $(HOST)/include/goofit/fitting/FitManagerMinuit1.h
RT_CONVERT += $(RT_ROOT_CONVERT)
RT_SIMPLE  += $(RT_ROOT_SIMPLE)
RT_BIN     := $(RT_CONVERT:%=examples/bin/reftests/convert/%) $(RT_SIMPLE:%=examples/bin/reftests/simple/%) $(RT_PDFS:%=examples/bin/reftests/PDFs/%)
RT_FLAGS   := $(HFLAGS) -Wno-unused-variable -Itests/ref_catch -I$(REF)/tests/include -I$(REF)/tests/convert -I$(REF)/tests/PDFs -I$(REF)/extern/fmt/include -DFMT_HEADER_ONLY
build/reftests/%.o: $(REF)/tests/%.cpp tests/ref_catch/catch2/catch_all.hpp $(wildcard $(HOST)/include/goofit/detail/*.h)
	@mkdir -p $(dir $@)
	$(CXX) $(RT_FLAGS) $(RT_EXTRA) -c $< -o $@
# one executable per test file, as upstream's goofit_catch_test() builds them (each with tests/catch_main.cpp)
examples/bin/reftests/%: build/reftests/%.o build/reftests/catch_main.o $(LIBDIR)/libgoofit_host.so
	@mkdir -p $(dir $@)
	$(CXX) $< build/reftests/catch_main.o $(LDSTD) -L$(LIBDIR) -lgoofit_host -lgoofit_b200 -lMinuit2 -Wl,-rpath,'$$ORIGIN/../../../../goofit_b200/lib' -o $@
.PRECIOUS: build/reftests/%.o
REFEX    += $(if $(wildcard $(REF)/tests/catch_main.cpp),$(RT_BIN),)
# The reference's ROOT-using examples, unmodified as well: GOOFIT_ROOT_FOUND=1 switches on the header-only ROOT conveniences of the
# drop-in headers (plotToROOT, make2D, setROOTStyle) and tests/ref_root stands in for ROOT itself (histograms in memory, canvases that
# draw nothing): neither machine has ROOT or a graphics stack. A user with ROOT passes its include path and libraries instead.
REFEX_ROOT := $(if $(wildcard $(REF)/examples/dalitz/dalitz.cpp),examples/bin/ref_dalitz examples/bin/ref_addition examples/bin/ref_simpleFit \
              examples/bin/ref_convolution examples/bin/ref_2d_plot examples/bin/ref_zachFit examples/bin/ref_SigGen examples/bin/ref_TDDP4 \
              examples/bin/ref_chisquare examples/bin/ref_exponential1 examples/bin/ref_exponential2,)
$(REFEX_ROOT): REFEX_EXTRA := -DGOOFIT_ROOT_FOUND=1 -Itests/ref_root -I$(REF)/extern/fmt/include -DFMT_HEADER_ONLY -Wno-sign-compare -Wno-unused-variable -Wno-unused-but-set-variable
# SigGen.cpp and TDDP4.cu name MCBooster's host containers (goofit_b200/host/include/mcbooster: the subset GenerateSig returns) and call
# thrust::count_if on them: Thrust's headers come with the CUDA toolkit, host backend
examples/bin/ref_SigGen examples/bin/ref_TDDP4: REFEX_EXTRA += -isystem /usr/local/cuda/include -DTHRUST_DEVICE_SYSTEM=THRUST_DEVICE_SYSTEM_CPP
$(REFEX_ROOT): $(wildcard tests/ref_root/*.h)
# exponential1.cpp fits through FitManagerMinuit1, chisquare.cpp drives a TMinuit of its own beside GooFit's fit (tests/ref_root/TMinuit.h, carried by Minuit2)
examples/bin/ref_exponential1 examples/bin/ref_exponential2: examples/bin/ref_exponential%: $(REF)/examples/exponential/exponential%.cpp $(LIBDIR)/libgoofit_host.so $(HOST)/include/goofit/fitting/FitManagerMinuit1.h
	@mkdir -p examples/bin
	$(CXX) $(HFLAGS) $(REFEX_EXTRA) -I$(REF)/extern/CLI11/include $< $(LDSTD) -L$(LIBDIR) -lgoofit_host -lgoofit_b200 -lMinuit2 -Wl,-rpath,'$$ORIGIN/../../goofit_b200/lib' -o $@
examples/bin/ref_TDDP4: $(REF)/examples/TDDP4/TDDP4.cu $(LIBDIR)/libgoofit_host.so $(HOST)/include/goofit/Application.h
	@mkdir -p examples/bin
	$(CXX) $(HFLAGS) $(REFEX_EXTRA) -I$(REF)/extern/CLI11/include -x c++ $< -x none $(LDSTD) -L$(LIBDIR) -lgoofit_host -lgoofit_b200 -lMinuit2 -Wl,-rpath,'$$ORIGIN/../../goofit_b200/lib' -o $@
REFEX    += $(REFEX_ROOT)
examples: examples/bin/exponential $(REFEX)
examples/bin/ref_%: $(REF)/examples/%/$$*.cpp $(LIBDIR)/libgoofit_host.so $(HOST)/include/goofit/Application.h
	@mkdir -p examples/bin
	$(CXX) $(HFLAGS) $(REFEX_EXTRA) -I$(REF)/extern/CLI11/include $< $(LDSTD) -L$(LIBDIR) -lgoofit_host -lgoofit_b200 -lMinuit2 -Wl,-rpath,'$$ORIGIN/../../goofit_b200/lib' -o $@
examples/bin/%: examples/%.cpp $(LIBDIR)/libgoofit_host.so
	@mkdir -p examples/bin
	$(CXX) $(HFLAGS) $< $(LDSTD) -L$(LIBDIR) -lgoofit_host -lgoofit_b200 -lMinuit2 -Wl,-rpath,'$$ORIGIN/../../goofit_b200/lib' -o $@

clean:
	rm -rf build $(LIBDIR)/*.so goofit_b200/_goofit*.so
.PHONY: all cuda host python examples clean
