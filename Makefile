# Build: product libraries (CUDA hot path + stand-alone host set-up) and the
# test-only oracle.  Everything is in-tree so the .so files travel with gpurun.
CXX       = /usr/bin/g++
NVCC     ?= /usr/local/cuda/bin/nvcc
CXXFLAGS  = -std=gnu++17 -O2 -fPIC -fno-fast-math -ffp-contract=off -Wall -Wno-unused-function
# -fmad=false: the stencils are HBM-bound, so contraction buys nothing, and without
# it every element-wise operator is bit-identical to the -ffp-contract=off oracle.
NVFLAGS   = -std=c++17 -O3 -gencode arch=compute_100a,code=sm_100a -lineinfo -Xcompiler -fPIC \
            -fmad=false -Xptxas -v

LIBDIR = gmcore_b200/lib
HOSTSRC = gmcore_b200/csrc/host_api.cpp
HOSTHDR = $(wildcard gmcore_b200/csrc/host_*.hpp) include/gmcore_host.h include/gmcore_gpu.h
GPUSRC  = gmcore_b200/csrc/gmcore_gpu.cu
GPUHDR  = $(wildcard gmcore_b200/csrc/*.cuh) include/gmcore_gpu.h
ORCSRC  = oracle/orc_api.cpp
ORCHDR  = $(wildcard oracle/*.hpp) include/gmcore_gpu.h

all: host gpu oracle

host: $(LIBDIR)/libgmcore_host.so
gpu: $(LIBDIR)/libgmcore_b200.so
oracle: oracle/_build/liboracle.so oracle/_build/liboracle_fast.so oracle/_build/libfastmath_host.so

$(LIBDIR)/libgmcore_host.so: $(HOSTSRC) $(HOSTHDR)
	@mkdir -p $(LIBDIR)
	$(CXX) $(CXXFLAGS) -shared -o $@ $(HOSTSRC) -lquadmath

$(LIBDIR)/libgmcore_b200.so: $(GPUSRC) $(GPUHDR)
	@mkdir -p $(LIBDIR)
	$(NVCC) $(NVFLAGS) -shared -o $@ $(GPUSRC) 2> $(LIBDIR)/ptxas.log || (cat $(LIBDIR)/ptxas.log; false)

# parity build (strict FP) and the timed CPU-baseline build (-O3 -march=native, OpenMP)
oracle/_build/liboracle.so: $(ORCSRC) $(ORCHDR)
	@mkdir -p oracle/_build
	$(CXX) $(CXXFLAGS) -shared -o $@ $(ORCSRC)
oracle/_build/liboracle_fast.so: $(ORCSRC) $(ORCHDR)
	@mkdir -p oracle/_build
	$(CXX) -std=c++17 -O3 -march=x86-64-v3 -fPIC -fno-fast-math -fopenmp -DORC_OPENMP -DORC_OPENMP_LOOPS -shared -o $@ $(ORCSRC)

# host build of the device's pow_pos / log_pos (tests only: accuracy against libquadmath, bit identity with the device)
oracle/_build/libfastmath_host.so: tests/fastmath_host.cpp gmcore_b200/csrc/fastmath.cuh
	@mkdir -p oracle/_build
	$(CXX) -std=gnu++17 -O2 -fPIC -fno-fast-math -ffp-contract=off -mfma -x c++ -shared -o $@ tests/fastmath_host.cpp -lquadmath

clean:
	rm -rf $(LIBDIR) oracle/_build

.PHONY: all host gpu oracle clean
