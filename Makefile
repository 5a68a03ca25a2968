# Build everything in-tree for sm_100a (B200). No GPU is needed to build.
#   make            -> build/libb200knn.so + executables under build/
#   make lib        -> only the shared library (what Python loads)
#   make sass       -> profiles/sass/*.sass listings of the hot kernels
NVCC      ?= /usr/local/cuda/bin/nvcc
CXX       ?= g++
ARCH      := -gencode arch=compute_100a,code=sm_100a
NVFLAGS   := $(ARCH) -lineinfo -O3 -std=c++17 -Xcompiler -fPIC,-ffp-contract=off -Xptxas -v \
             --fmad=false -prec-sqrt=true -prec-div=true -Iinclude -Icsrc
CXXFLAGS  := -O2 -std=c++17 -fPIC -ffp-contract=off -mfma -pthread -Iinclude -Icsrc
BUILD     := build

CU_SRCS   := csrc/radix_sort.cu csrc/build.cu csrc/query.cu csrc/pv.cu csrc/redist.cu csrc/api.cu csrc/dist.cu csrc/noalloc.cu csrc/neighbors.cu csrc/verify.cu
# tcgen05 / TMA code: GPU library only (the host emulation below has no tensor cores to emulate)
TC_SRCS   := csrc/brute_tile.cu
CU_OBJS   := $(patsubst csrc/%.cu,$(BUILD)/%.o,$(CU_SRCS) $(TC_SRCS))
CPP_OBJS  := $(BUILD)/cpu_engine.o

LIB       := $(BUILD)/libb200knn.so

all: lib

lib: $(LIB)

$(BUILD)/%.o: csrc/%.cu csrc/*.h csrc/*.cuh include/b200knn.h
	@mkdir -p $(BUILD)
	$(NVCC) $(NVFLAGS) -c $< -o $@ 2> $(BUILD)/$*.ptxas.log || (cat $(BUILD)/$*.ptxas.log; false)

$(BUILD)/%.o: csrc/%.cpp csrc/*.h include/b200knn.h
	@mkdir -p $(BUILD)
	$(CXX) $(CXXFLAGS) -c $< -o $@

$(LIB): $(CU_OBJS) $(CPP_OBJS)
	$(NVCC) $(ARCH) -shared -o $@ $^ -cudart static -lpthread -ldl

clean:
	rm -rf $(BUILD)

.PHONY: all lib clean

# ---- executables (same names and command lines as the reference's) ----
APPS := $(BUILD)/knnrun $(BUILD)/cudaMpiKNN_unorderedData $(BUILD)/cudaMpiKNN_prePartitionedData
APPFLAGS := -O2 -std=c++17 -Iinclude -Xcompiler -ffp-contract=off

apps: $(APPS)
all: apps

$(BUILD)/knnrun: apps/knnrun.cpp
	@mkdir -p $(BUILD)
	$(CXX) -O2 -std=c++17 $< -o $@

$(BUILD)/backend.o: apps/backend.cpp apps/knn_apps.hpp include/b200knn.h
	$(NVCC) $(APPFLAGS) -x cu $(ARCH) -c $< -o $@

$(BUILD)/cudaMpiKNN_unorderedData: apps/unordered.cpp $(BUILD)/backend.o $(LIB) apps/knn_apps.hpp
	$(NVCC) $(APPFLAGS) $(ARCH) $< $(BUILD)/backend.o -o $@ -L$(BUILD) -lb200knn -Xlinker -rpath,'$$ORIGIN'

$(BUILD)/cudaMpiKNN_prePartitionedData: apps/prepartitioned.cpp $(BUILD)/backend.o $(LIB) apps/knn_apps.hpp
	$(NVCC) $(APPFLAGS) $(ARCH) $< $(BUILD)/backend.o -o $@ -L$(BUILD) -lb200knn -Xlinker -rpath,'$$ORIGIN'

.PHONY: apps

# ---- examples ----
examples: $(BUILD)/sample_knn
$(BUILD)/sample_knn: examples/sample_knn.c $(LIB) include/b200knn.h
	gcc -std=c99 -O2 -Iinclude -I/usr/local/cuda/include $< -o $@ -L$(BUILD) -lb200knn \
	    -L/usr/local/cuda/lib64 -lcudart -lm -Wl,-rpath,'$$ORIGIN' -Wl,-rpath,/usr/local/cuda/lib64
.PHONY: examples

# ---- SASS listings of every kernel (profiles/sass/, no GPU needed) ----
sass: lib
	python tools/dump_sass.py
.PHONY: sass

# ---- host emulation of the kernels (tests/simt): the same sources compiled as C++, threads = fibers ----
SIMT_DIR  := $(BUILD)/simt
SIMT_SRCS := $(patsubst csrc/%.cu,$(SIMT_DIR)/%.cpp,$(CU_SRCS))
SIMT_LIB  := $(BUILD)/libb200knn_simt.so
SIMT_FLAGS := -O1 -g -std=c++17 -fPIC -ffp-contract=off -mfma -pthread -fpermissive -w -Itests/simt -Iinclude -Icsrc

simt: $(SIMT_LIB)
$(SIMT_DIR)/%.cpp: csrc/%.cu tests/simt/preprocess.py
	@mkdir -p $(SIMT_DIR)
	python tests/simt/preprocess.py $< $@
$(SIMT_DIR)/%.o: $(SIMT_DIR)/%.cpp tests/simt/simt_host.h csrc/*.h csrc/*.cuh include/b200knn.h
	$(CXX) $(SIMT_FLAGS) -c $< -o $@
$(SIMT_DIR)/simt_runtime.o: tests/simt/simt_runtime.cpp tests/simt/simt_host.h
	@mkdir -p $(SIMT_DIR)
	$(CXX) $(SIMT_FLAGS) -c $< -o $@
$(SIMT_LIB): $(patsubst %.cpp,%.o,$(SIMT_SRCS)) $(SIMT_DIR)/simt_runtime.o $(CPP_OBJS)
	$(CXX) -shared -o $@ $^ -lpthread -ldl
.PHONY: simt
.SECONDARY: $(SIMT_SRCS)

# ---- the same emulation under ThreadSanitizer: race detection for the protocols BETWEEN ranks ----
#   make simt-tsan && LD_PRELOAD=$$(gcc -print-file-name=libtsan.so) B200KNN_SIMT_LIB=build/libb200knn_simt_tsan.so \
#       python tests/simt/multirank_worker.py 2 uniform chunks
TSAN_DIR := $(BUILD)/simt_tsan
TSAN_LIB := $(BUILD)/libb200knn_simt_tsan.so
TSAN_CXX ?= /usr/bin/g++
TSAN_FLAGS := $(SIMT_FLAGS) -fsanitize=thread --param tsan-instrument-func-entry-exit=0
simt-tsan: $(TSAN_LIB)
$(TSAN_DIR)/%.o: $(SIMT_DIR)/%.cpp tests/simt/simt_host.h csrc/*.h csrc/*.cuh include/b200knn.h
	@mkdir -p $(TSAN_DIR)
	$(TSAN_CXX) $(TSAN_FLAGS) -c $< -o $@
$(TSAN_DIR)/simt_runtime.o: tests/simt/simt_runtime.cpp tests/simt/simt_host.h
	@mkdir -p $(TSAN_DIR)
	$(TSAN_CXX) $(TSAN_FLAGS) -c $< -o $@
$(TSAN_DIR)/cpu_engine.o: csrc/cpu_engine.cpp csrc/*.h include/b200knn.h
	@mkdir -p $(TSAN_DIR)
	$(TSAN_CXX) $(CXXFLAGS) -fsanitize=thread --param tsan-instrument-func-entry-exit=0 -c $< -o $@
$(TSAN_LIB): $(patsubst $(SIMT_DIR)/%.cpp,$(TSAN_DIR)/%.o,$(SIMT_SRCS)) $(TSAN_DIR)/simt_runtime.o $(TSAN_DIR)/cpu_engine.o
	$(TSAN_CXX) -shared -fsanitize=thread -o $@ $^ -lpthread -ldl
.PHONY: simt-tsan

# ---- the same emulation under AddressSanitizer: out-of-bounds / use-after-free in "device" memory ----
#   make simt-asan && LD_PRELOAD=$$(gcc -print-file-name=libasan.so) ASAN_OPTIONS=detect_leaks=0 \
#       B200KNN_SIMT_LIB=build/libb200knn_simt_asan.so python tests/simt/multirank_worker.py 2 uniform chunks
ASAN_DIR := $(BUILD)/simt_asan
ASAN_LIB := $(BUILD)/libb200knn_simt_asan.so
ASAN_FLAGS := $(SIMT_FLAGS) -fsanitize=address -fno-omit-frame-pointer
simt-asan: $(ASAN_LIB)
$(ASAN_DIR)/%.o: $(SIMT_DIR)/%.cpp tests/simt/simt_host.h csrc/*.h csrc/*.cuh include/b200knn.h
	@mkdir -p $(ASAN_DIR)
	$(TSAN_CXX) $(ASAN_FLAGS) -c $< -o $@
$(ASAN_DIR)/simt_runtime.o: tests/simt/simt_runtime.cpp tests/simt/simt_host.h
	@mkdir -p $(ASAN_DIR)
	$(TSAN_CXX) $(ASAN_FLAGS) -c $< -o $@
$(ASAN_DIR)/cpu_engine.o: csrc/cpu_engine.cpp csrc/*.h include/b200knn.h
	@mkdir -p $(ASAN_DIR)
	$(TSAN_CXX) $(CXXFLAGS) -fsanitize=address -c $< -o $@
$(ASAN_LIB): $(patsubst $(SIMT_DIR)/%.cpp,$(ASAN_DIR)/%.o,$(SIMT_SRCS)) $(ASAN_DIR)/simt_runtime.o $(ASAN_DIR)/cpu_engine.o
	$(TSAN_CXX) -shared -fsanitize=address -o $@ $^ -lpthread -ldl
.PHONY: simt-asan
