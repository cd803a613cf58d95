# Plain-make build for C/C++ users (the Python side uses cuda_mpi_parallel_b200/utils/build.py with the same flags).
# The reference's build recipe is one line in a comment: nvcc CUDACG.cu -lcublas -lcusparse (CUDACG.cu:1).
#
#   make            libcgb200.so + bin/cgb200
#   make examples   bin/poisson3d (plain C on the C ABI)
#   make baselines  libcgb200_baseline.so (B3) + bin/refstyle_b4 (B4; needs the system libnccl)
#   make OUT=/tmp/x out-of-tree objects and products
NVCC    ?= nvcc
CXX     ?= g++
OUT     ?= .
OBJDIR  := $(OUT)/build/make
LIBDIR  := $(OUT)/cuda_mpi_parallel_b200/lib
BINDIR  := $(OUT)/bin
ARCH    := -gencode arch=compute_100a,code=sm_100a
NVFLAGS := $(ARCH) -O3 -std=c++17 -lineinfo -Xcompiler -fPIC,-Wall,-Wno-unused-function -Iinclude -Icsrc

SRCS := $(shell find csrc -name '*.cu' -o -name '*.cpp' | sort)
HDRS := $(shell find csrc include -name '*.hpp' -o -name '*.cuh' -o -name '*.h')
OBJS := $(patsubst %,$(OBJDIR)/%.o,$(SRCS))
BSRC := $(filter-out %_main.cu,$(wildcard baseline/src/*.cu))

all: $(LIBDIR)/libcgb200.so $(BINDIR)/cgb200

$(OBJDIR)/%.o: % $(HDRS)
	@mkdir -p $(dir $@)
	$(NVCC) $(NVFLAGS) -c $< -o $@

$(LIBDIR)/libcgb200.so: $(OBJS)
	@mkdir -p $(LIBDIR)
	$(NVCC) $(ARCH) -shared -o $@ $(OBJS)

$(BINDIR)/cgb200: tools/cgb200_main.cpp include/cgb200.h $(LIBDIR)/libcgb200.so
	@mkdir -p $(BINDIR)
	$(CXX) -O2 -std=c++17 -Wall -Iinclude $< -o $@ -L$(LIBDIR) -lcgb200 '-Wl,-rpath,$$ORIGIN/../cuda_mpi_parallel_b200/lib'

baselines: $(LIBDIR)/libcgb200_baseline.so $(BINDIR)/refstyle_b4

$(LIBDIR)/libcgb200_baseline.so: $(BSRC) $(HDRS)
	@mkdir -p $(LIBDIR)
	$(NVCC) $(NVFLAGS) -shared -o $@ $(BSRC) -lcusparse -lcublas

$(BINDIR)/refstyle_b4: baseline/src/ref_b4_main.cu include/cgb200.h $(LIBDIR)/libcgb200.so
	@mkdir -p $(BINDIR)
	$(NVCC) $(ARCH) -O3 -std=c++17 -lineinfo -Iinclude $< -o $@ -L$(LIBDIR) -lcgb200 -lnccl -lcusparse -lcublas -lpthread \
	    -Xlinker -rpath -Xlinker '$$ORIGIN/../cuda_mpi_parallel_b200/lib'

examples: $(BINDIR)/poisson3d

$(BINDIR)/poisson3d: examples/poisson3d.c include/cgb200.h $(LIBDIR)/libcgb200.so
	@mkdir -p $(BINDIR)
	$(CC) -std=c99 -O2 -Wall -Iinclude $< -o $@ -L$(LIBDIR) -lcgb200 '-Wl,-rpath,$$ORIGIN/../cuda_mpi_parallel_b200/lib' -lm

clean:
	rm -rf $(OBJDIR)

.PHONY: all baselines examples clean
