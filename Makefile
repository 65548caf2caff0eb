# Builds libpopsift_b200.so (C ABI + the C++ drop-in classes) for sm_100a, in-tree.
NVCC   ?= nvcc
ARCH   := -gencode arch=compute_100a,code=sm_100a
NVFLAGS := -std=c++17 -O3 $(ARCH) -lineinfo -Xcompiler -fPIC -Xcompiler -ffp-contract=off -Iinclude
CSRC   := popsift_b200/csrc
OBJDIR := build/obj
LIB    := popsift_b200/libpopsift_b200.so

CU_SRCS  := $(CSRC)/k_pyramid.cu $(CSRC)/k_extrema.cu $(CSRC)/k_filter.cu $(CSRC)/k_orient.cu $(CSRC)/k_desc.cu $(CSRC)/k_match.cu $(CSRC)/pipeline.cu
CPP_SRCS := $(CSRC)/tables.cpp $(wildcard $(CSRC)/host/*.cpp)
OBJS := $(patsubst $(CSRC)/%.cu,$(OBJDIR)/%.o,$(CU_SRCS)) $(patsubst $(CSRC)/%.cpp,$(OBJDIR)/%.o,$(CPP_SRCS))

all: $(LIB) tests/cpp/api_demo tests/cpp/match_demo popsift_b200/popsift-demo

# the reference's command-line front end (src/application/main.cpp), Boost-free
popsift_b200/popsift-demo: apps/popsift_demo.cpp $(LIB) $(wildcard include/popsift/*.h)
	g++ -std=c++17 -O2 -Iinclude -o $@ $< -Lpopsift_b200 -lpopsift_b200 -Wl,-rpath,'$$ORIGIN' -lpthread

tests/cpp/%_demo: tests/cpp/%_demo.cpp $(LIB) $(wildcard include/popsift/*.h)
	g++ -std=c++17 -O2 -Iinclude -o $@ $< -Lpopsift_b200 -lpopsift_b200 -Wl,-rpath,'$$ORIGIN/../../popsift_b200' -lpthread

$(OBJDIR)/%.o: $(CSRC)/%.cu $(CSRC)/common.h $(CSRC)/kernels.h include/psift.h
	@mkdir -p $(dir $@)
	$(NVCC) $(NVFLAGS) -c -o $@ $<

$(OBJDIR)/%.o: $(CSRC)/%.cpp $(CSRC)/common.h include/psift.h
	@mkdir -p $(dir $@)
	$(NVCC) $(NVFLAGS) -c -o $@ $<

$(LIB): $(OBJS)
	$(NVCC) $(ARCH) -shared -o $@ $^ -lcudart

clean:
	rm -rf build $(LIB)
.PHONY: all clean
