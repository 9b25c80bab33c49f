# Builds the product library (CUDA backend + C++ host layer) and, separately, the test oracle.
#   make            -> raysnail_b200/libraysnail_b200.so   (sm_100a only)
#   make oracle     -> oracle/liboracle.so                 (test infrastructure)
NVCC      ?= /usr/local/cuda/bin/nvcc
CXX       ?= g++
ARCH      := -gencode arch=compute_100a,code=sm_100a
# TUNE: extra -D flags for kernel tuning experiments; VARIANT: suffix for an experimental build
# (make VARIANT=_t1 TUNE="-DRS_TUNE_SHADE_LEAN=3" builds raysnail_b200/libraysnail_b200_t1.so,
#  selected at run time with RAYSNAIL_B200_LIB=<path>)
TUNE      ?=
VARIANT   ?=
NVFLAGS   := -std=c++17 -O3 -lineinfo $(ARCH) -Xcompiler -fPIC,-O3 -Xptxas -v --expt-relaxed-constexpr $(TUNE)
# -ffp-contract=off: Rust never fuses a*b+c on its own, so neither does the host layer
CXXFLAGS  := -std=c++17 -O2 -fPIC -Wall -Wextra -march=x86-64-v3 -ffp-contract=off
BUILD     := build$(VARIANT)
CSRC      := raysnail_b200/csrc
LIB       := raysnail_b200/libraysnail_b200$(VARIANT).so

HOST_SRCS := $(CSRC)/host/scene.cpp $(CSRC)/host/sdl_parser.cpp $(CSRC)/host/recipes.cpp \
             $(CSRC)/host/image_io.cpp $(CSRC)/host/host_api.cpp
HOST_OBJS := $(patsubst $(CSRC)/host/%.cpp,$(BUILD)/host_%.o,$(HOST_SRCS))
CU_OBJS   := $(BUILD)/kernels.o $(BUILD)/backend.o
DEV_HDRS  := $(CSRC)/device/rt_math.cuh $(CSRC)/device/rt_scene.cuh $(CSRC)/device/kernels.cuh include/raysnail_b200.h

all: $(LIB)

$(BUILD):
	mkdir -p $(BUILD)

$(BUILD)/host_%.o: $(CSRC)/host/%.cpp $(CSRC)/host/scene.hpp $(CSRC)/host/seed_random.hpp include/raysnail_b200.h include/raysnail_b200_host.h | $(BUILD)
	$(CXX) $(CXXFLAGS) -c $< -o $@

$(BUILD)/kernels.o: $(CSRC)/device/kernels.cu $(DEV_HDRS) | $(BUILD)
	$(NVCC) $(NVFLAGS) -c $< -o $@ 2> $(BUILD)/kernels.ptxas.log || (cat $(BUILD)/kernels.ptxas.log; false)

$(BUILD)/backend.o: $(CSRC)/backend.cu $(DEV_HDRS) | $(BUILD)
	$(NVCC) $(NVFLAGS) -c $< -o $@

$(LIB): $(HOST_OBJS) $(CU_OBJS)
	$(NVCC) -shared $(ARCH) -o $@ $^ -cudart static -lz

oracle: oracle/liboracle.so

# strict IEEE: no FMA contraction beyond the explicit fma() calls that restate mul_add
oracle/liboracle.so: oracle/oracle.cpp include/raysnail_b200.h
	$(CXX) -std=c++17 -O3 -march=x86-64-v3 -ffp-contract=off -fPIC -shared -pthread -Wall -Wextra $< -o $@

clean:
	rm -rf $(BUILD) $(LIB) oracle/liboracle.so

.PHONY: all oracle clean
