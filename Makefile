# b200reduce — build for NVIDIA B200 (sm_100a) only.
# Replaces the reference's SDK build (cuda/C/common/common.mk: -gencode sm_10/sm_13/sm_20, -lcutil -lshrutil) and
# mpi/Makefile (mpixlc / mpicc). Make variables kept in the spirit of common.mk:156-176,339-363:
#   dbg=1 (-G -O0)   ptxas=1 (-Xptxas -v)   keep=1 (-keep)   verbose=1
NVCC      ?= /usr/local/cuda/bin/nvcc
CXX       ?= g++
ARCH      := -gencode arch=compute_100a,code=sm_100a
PKG       := cuda_mpi_reductions_b200
LIBDIR    := $(PKG)/lib
BINDIR    := bin
OBJDIR    := build

NVCCFLAGS := $(ARCH) -std=c++17 -lineinfo -Xcompiler -fPIC,-Wall,-Wno-unused-function -Iinclude -Isrc --expt-relaxed-constexpr
CXXFLAGS  := -std=c++17 -O2 -fPIC -Wall -Iinclude -Isrc -pthread
ifeq ($(dbg),1)
  NVCCFLAGS += -G -O0
else
  NVCCFLAGS += -O3
endif
ifeq ($(ptxas),1)
  NVCCFLAGS += -Xptxas -v
endif
ifeq ($(keep),1)
  NVCCFLAGS += -keep
endif
ifneq ($(verbose),1)
  Q := @
endif

CU_SRCS  := $(wildcard src/core/*.cu) $(wildcard src/comm/*.cu) $(wildcard src/coll/*.cu)
CPP_SRCS := $(wildcard src/core/*.cpp) $(wildcard src/comm/*.cpp) $(wildcard src/coll/*.cpp) $(wildcard src/util/*.cpp)
CU_OBJS  := $(patsubst src/%.cu,$(OBJDIR)/%.cu.o,$(CU_SRCS))
CPP_OBJS := $(patsubst src/%.cpp,$(OBJDIR)/%.cpp.o,$(CPP_SRCS))
HDRS     := $(wildcard include/*.h) $(wildcard src/*/*.h) $(wildcard src/*/*.cuh) $(wildcard src/*/*.inl)

LIB      := $(LIBDIR)/libb2r.so
BENCHES  := $(BINDIR)/local_sweep $(BINDIR)/bench_reduce $(BINDIR)/bench_collective $(BINDIR)/launch_gap

.PHONY: all lib bins minimpi baseline examples clean sass
all: lib bins minimpi

lib: $(LIB)

$(OBJDIR)/%.cu.o: src/%.cu $(HDRS)
	@mkdir -p $(dir $@)
	@echo "NVCC $<"
	$(Q)$(NVCC) $(NVCCFLAGS) -c $< -o $@

$(OBJDIR)/%.cpp.o: src/%.cpp $(HDRS)
	@mkdir -p $(dir $@)
	@echo "CXX  $<"
	$(Q)$(CXX) $(CXXFLAGS) -I/usr/local/cuda/include -c $< -o $@

$(LIB): $(CU_OBJS) $(CPP_OBJS)
	@mkdir -p $(LIBDIR)
	@echo "LINK $@"
	$(Q)$(NVCC) $(ARCH) -shared -cudart static -o $@ $^ -lpthread -ldl -lrt

bins: $(BENCHES)

$(BINDIR)/%: bench/%.cu $(LIB) $(HDRS) $(wildcard bench/*.h)
	@mkdir -p $(BINDIR)
	@echo "NVCC $<"
	$(Q)$(NVCC) $(NVCCFLAGS) -o $@ $< -L$(LIBDIR) -lb2r -ldl -Xlinker -rpath,'$$ORIGIN/../$(LIBDIR)'

minimpi:
	$(Q)$(MAKE) -C minimpi $(if $(Q),-s,)

baseline:
	$(Q)bash baseline/build_baselines.sh

# MPI-launched GPU example (one mini-MPI rank per GPU; MPI carries only rank/size/bootstrap id)
examples: lib minimpi
	@mkdir -p $(BINDIR)
	$(Q)$(NVCC) $(NVCCFLAGS) -Iminimpi -o $(BINDIR)/reduce_gpu_mpi examples/reduce_gpu_mpi.cu minimpi/minimpi.o -L$(LIBDIR) -lb2r -lrt -Xlinker -rpath,'$$ORIGIN/../$(LIBDIR)'

sass: $(LIB)
	@mkdir -p profiles/sass
	cuobjdump -sass $(LIB) > profiles/sass/libb2r.sass

clean:
	rm -rf $(OBJDIR) $(BINDIR) $(LIB)
	-$(MAKE) -C minimpi clean
