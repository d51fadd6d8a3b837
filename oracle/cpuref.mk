# The REFERENCE's whole CPU backend compiled from its own sources where they lie under $(REF)
# (TEST INFRASTRUCTURE: the checker for per-op and per-model parity; never linked by mllm_b200/).
# The reference's build system (cmake) is not run: this is g++ on the source files the reference's
# CMakeLists glob (CMakeLists.txt:186-193, mllm/backends/cpu/CMakeLists.txt:2-8), with its x86 flags
# (-mavx2 -mf16c, CMakeLists.txt:174-175) minus -march=native so the .so also runs on the GPU box's host.
# Output only under oracle/_ref/ (git-ignored; travels to the GPU box).
REF    ?= /root/reference
CXX    := /usr/bin/g++
HERE   := $(dir $(abspath $(lastword $(MAKEFILE_LIST))))
OUT    := $(HERE)_ref
OBJ    := $(OUT)/cpuobj
CPU    := $(REF)/mllm/backends/cpu
SRCS   := $(wildcard $(REF)/mllm/*.cpp) $(wildcard $(REF)/mllm/memory/*.cpp) $(wildcard $(REF)/mllm/processor/*.cpp) \
          $(wildcard $(REF)/mllm/tokenizers/*.cpp) $(wildcard $(REF)/mllm/tokenizers/*/*.cpp) \
          $(wildcard $(CPU)/*.cpp) $(wildcard $(CPU)/compute/*.cpp) $(wildcard $(CPU)/third_party/ggml/*.cpp) $(wildcard $(CPU)/op/*.cpp)
OBJS   := $(patsubst $(REF)/%.cpp,$(OBJ)/%.o,$(filter-out %/AudioProcess.cpp,$(SRCS)))
INC    := -I$(REF)/mllm -I$(REF)/third_party -I$(REF)/third_party/fmt/include -I$(CPU) -I$(CPU)/compute -I$(CPU)/third_party/ggml -I$(REF)/include -I$(REF)
FLAGS  := -std=c++17 -O2 -mavx2 -mf16c -mfma -fopenmp -fPIC -w -DNDEBUG

all: $(OUT)/libmllm_cpu_ref.so

$(OBJ)/%.o: $(REF)/%.cpp
	@mkdir -p $(dir $@)
	$(CXX) $(FLAGS) $(INC) -c $< -o $@

$(OUT)/libmllm_cpu_ref.so: $(OBJS) $(HERE)ref_ops_shim.cpp
	$(CXX) $(FLAGS) $(INC) -shared -Wl,-z,defs -o $@ $(HERE)ref_ops_shim.cpp $(REF)/third_party/fmt/src/format.cc $(OBJS) -lm

.PHONY: all
