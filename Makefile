# Build everything in-tree (the built .so/binaries travel to the GPU box with the gpurun snapshot).
#   make            -> libgfd_b200.so, gpu-feature-discovery, fake NVML, oracle C helpers
#   make asan       -> ASan/UBSan build of the host-only parts under build/asan (tests use it when present)
NVCC      ?= /usr/local/cuda/bin/nvcc
CXX       ?= g++
CC        ?= gcc
CUDA_INC  ?= /usr/local/cuda/include
PKG       := gpu-feature-discovery_b200
CSRC      := $(PKG)/csrc
GENCODE   := -gencode arch=compute_100a,code=sm_100a
NVCCFLAGS := $(GENCODE) -O3 -lineinfo -std=c++17 -Iinclude -I$(CSRC) --fmad=false \
             -Xcompiler -fPIC,-fvisibility=hidden,-Wall,-Wextra,-ffp-contract=off -Xptxas -v
CXXFLAGS  := -O2 -std=c++17 -Wall -Wextra -fPIC -Iinclude -ffp-contract=off -ffunction-sections -fdata-sections

LIB       := $(PKG)/libgfd_b200.so
BIN       := $(PKG)/gpu-feature-discovery
FAKE      := $(PKG)/fake_nvml/lib/libnvidia-ml.so.1
FAKE_NOMIG:= $(PKG)/fake_nvml/lib_nomig/libnvidia-ml.so.1
FAKE_CUDA := $(PKG)/fake_cuda/lib/libcuda.so.1
ORACLE_SO := oracle/libgfd_oracle_c.so
REPLAY    := oracle/ref_replay

LIB_SRCS  := $(CSRC)/gfd_nvml.cpp $(CSRC)/gfd_mig_profile.cpp $(CSRC)/gfd_cuda_mgr.cpp $(CSRC)/gfd_probe.cu
LABELER_SRCS := $(wildcard $(PKG)/labeler/*.cpp)

TARGETS := $(LIB) $(FAKE) $(FAKE_NOMIG) $(FAKE_CUDA) $(ORACLE_SO)
ifneq ($(wildcard oracle/ref_replay.c),)
TARGETS += $(REPLAY)
endif
ifneq ($(LABELER_SRCS),)
TARGETS += $(BIN)
endif
DEMO := examples/cabi_demo
TARGETS += $(DEMO)
YAML_DUMP := tools/yaml_dump
TARGETS += $(YAML_DUMP)

all: $(TARGETS)

$(LIB): $(LIB_SRCS) $(CSRC)/gfd_internal.h include/gfd_b200.h
	@mkdir -p build
	$(NVCC) $(NVCCFLAGS) -shared -cudart static -x cu $(LIB_SRCS) -o $@ -ldl -lpthread 2> build/ptxas_$(notdir $@).log || (cat build/ptxas_$(notdir $@).log; false)
	@grep -E "registers|spill" build/ptxas_$(notdir $@).log | sort | uniq -c | head -20

$(FAKE): $(PKG)/fake_nvml/fake_nvml.c
	@mkdir -p $(dir $@)
	$(CC) -O2 -Wall -Wextra -fPIC -shared -fvisibility=hidden -Wl,-soname,libnvidia-ml.so.1 -I$(CUDA_INC) -o $@ $< -lpthread

$(FAKE_NOMIG): $(PKG)/fake_nvml/fake_nvml.c
	@mkdir -p $(dir $@)
	$(CC) -O2 -Wall -Wextra -fPIC -shared -fvisibility=hidden -Wl,-soname,libnvidia-ml.so.1 -DFAKE_NO_MIG_SYMBOLS -I$(CUDA_INC) -o $@ $< -lpthread

$(FAKE_CUDA): $(PKG)/fake_cuda/fake_cuda.c
	@mkdir -p $(dir $@)
	$(CC) -O2 -Wall -Wextra -fPIC -shared -fvisibility=hidden -Wl,-soname,libcuda.so.1 -o $@ $<

$(ORACLE_SO): oracle/checksum_oracle.c
	$(CC) -O2 -Wall -Wextra -fPIC -shared -o $@ $<

$(REPLAY): oracle/ref_replay.c
	$(CC) -O2 -Wall -Wextra -I$(CUDA_INC) -o $@ $< -ldl -lm

# The labeler links the HOST-ONLY translation units of the C ABI in statically (same sources as the .so) and dlopen()s
# libgfd_b200.so only for the opt-in probes (labeler/probe_loader.cpp): no CUDA start-up cost on the default pass.
CORE_SRCS := $(CSRC)/gfd_nvml.cpp $(CSRC)/gfd_mig_profile.cpp $(CSRC)/gfd_cuda_mgr.cpp
$(BIN): $(LABELER_SRCS) $(CORE_SRCS) $(CSRC)/gfd_internal.h $(LIB) include/gfd_b200.h $(wildcard $(PKG)/labeler/*.hpp)
	$(CXX) $(CXXFLAGS) -fPIE -I$(CSRC) -I$(CUDA_INC) $(LABELER_SRCS) $(CORE_SRCS) -o $@ -static-libstdc++ -static-libgcc -Wl,--gc-sections -s -ldl -lpthread

$(DEMO): examples/cabi_demo.c $(LIB) include/gfd_b200.h
	$(CC) -O2 -Wall -Wextra -Iinclude $< -o $@ -L$(PKG) -lgfd_b200 -Wl,-rpath,'$$ORIGIN/../$(PKG)'

# test helper: the labeler's YAML reader as a filter (tests/test_yaml_reader.py compares it with PyYAML)
$(YAML_DUMP): tools/yaml_dump.cpp $(PKG)/labeler/yaml_lite.cpp $(PKG)/labeler/yaml_lite.hpp
	$(CXX) -O1 -std=c++17 -Wall -Wextra $< $(PKG)/labeler/yaml_lite.cpp -o $@

clean:
	rm -f $(LIB) $(BIN) $(FAKE) $(FAKE_NOMIG) $(FAKE_CUDA) $(ORACLE_SO) $(REPLAY)

.PHONY: all clean

# ---- sanitizer build of everything host-side: `make asan`, then
#      GFD_TEST_BINDIR=build/asan python -m pytest tests/test_cli_parity.py tests/test_cli_flags_and_lifecycle.py
ASAN_DIR := build/asan
ASAN_FLAGS := -fsanitize=address,undefined -fno-omit-frame-pointer -g -O1
asan:
	@mkdir -p $(ASAN_DIR)
	$(NVCC) $(GENCODE) -O1 -g -lineinfo -std=c++17 -Iinclude -I$(CSRC) --fmad=false \
	    -Xcompiler -fPIC,-fvisibility=hidden,-Wall,-Wextra,-ffp-contract=off,-fsanitize=address,-fsanitize=undefined,-fno-omit-frame-pointer \
	    -shared -cudart static -x cu $(LIB_SRCS) -o $(ASAN_DIR)/libgfd_b200.so -ldl -lpthread -lasan -lubsan
	/usr/bin/g++ $(CXXFLAGS) $(ASAN_FLAGS) -fPIE -I$(CSRC) -I$(CUDA_INC) $(LABELER_SRCS) $(CORE_SRCS) -o $(ASAN_DIR)/gpu-feature-discovery -ldl -lpthread
.PHONY: asan

# ThreadSanitizer build of the labeler (the per-GPU attach/query threads): `make tsan`, then
#      GFD_TEST_BINDIR=build/tsan python -m pytest tests/test_cli_parity.py
tsan:
	@mkdir -p build/tsan
	/usr/bin/g++ -O1 -g -std=c++17 -fsanitize=thread -fPIE -Iinclude -I$(CSRC) -I$(CUDA_INC) $(LABELER_SRCS) $(CORE_SRCS) -o build/tsan/gpu-feature-discovery -ldl -lpthread
.PHONY: tsan
