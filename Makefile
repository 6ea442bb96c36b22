# Builds the product library (CUDA, sm_100a), the host tools and the test-only CPU oracle.
#   make lib      -> bucket-map_b200/libbucketmap_b200.so   (C-ABI of include/bucketmap_b200.h)
#   make tools    -> bucket-map_b200/bin/bucketmap, bucketmap_align   (C++ host over the C-ABI), bucketmap_analyze
#   make oracle   -> oracle/libbm_oracle.so                 (test infrastructure only)
#   make ref      -> oracle/_ref/...                        (reference headers built against oracle/shim; needs /root/reference)
NVCC      ?= nvcc
# the environment may point CXX at a wrapper whose libstdc++.so link is dangling (static libstdc++ in a .so crashes
# inside a Python process that already holds the shared one): always use the g++ on PATH
CXX       := g++
PKG       := bucket-map_b200
CSRC      := $(PKG)/csrc
BUILD     := build
CUFLAGS   := -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -std=c++17 -Xcompiler -fPIC,-Wall
CU_SRCS   := $(wildcard $(CSRC)/*.cu)
CU_OBJS   := $(patsubst $(CSRC)/%.cu,$(BUILD)/%.o,$(CU_SRCS))
LIB       := $(PKG)/libbucketmap_b200.so

all: lib oracle tools

lib: $(LIB)

$(BUILD)/%.o: $(CSRC)/%.cu $(CSRC)/bm_internal.cuh include/bucketmap_b200.h
	@mkdir -p $(BUILD)
	$(NVCC) $(CUFLAGS) -c $< -o $@

$(LIB): $(CU_OBJS)
	$(NVCC) -shared -o $@ $(CU_OBJS) -lcudart

oracle: oracle/libbm_oracle.so

oracle/libbm_oracle.so: oracle/bm_oracle.cpp oracle/bm_oracle.h
	$(CXX) -O3 -std=c++17 -fPIC -shared -pthread -o $@ oracle/bm_oracle.cpp

HOST_SRCS := $(wildcard $(PKG)/host/*.cpp)
HOST_HDRS := $(wildcard $(PKG)/host/*.hpp) include/bucketmap_b200.h

tools: $(PKG)/bin/bucketmap $(PKG)/bin/bucketmap_align $(PKG)/bin/bucketmap_analyze

# accuracy metrics of a SAM file against a ground truth (host-only evaluation tool)
$(PKG)/bin/bucketmap_analyze: $(PKG)/tools/sam_analyze.cpp
	@mkdir -p $(PKG)/bin
	$(CXX) -O2 -std=c++17 -Wall -o $@ $<

$(PKG)/bin/bucketmap: $(HOST_SRCS) $(HOST_HDRS) $(LIB)
	@mkdir -p $(PKG)/bin
	$(CXX) -O2 -std=c++17 -Iinclude -o $@ $(HOST_SRCS) -L$(PKG) -lbucketmap_b200 -Wl,-rpath,'$$ORIGIN/..' -pthread

$(PKG)/bin/bucketmap_align: $(HOST_SRCS) $(HOST_HDRS) $(LIB)
	@mkdir -p $(PKG)/bin
	$(CXX) -O2 -std=c++17 -DBM_ALIGN -Iinclude -o $@ $(HOST_SRCS) -L$(PKG) -lbucketmap_b200 -Wl,-rpath,'$$ORIGIN/..' -pthread

ref:
	$(MAKE) -C oracle ref

clean:
	rm -rf $(BUILD) $(LIB) oracle/libbm_oracle.so $(PKG)/bin

.PHONY: all lib oracle tools ref clean
