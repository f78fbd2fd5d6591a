# Builds libplsb.so (the sm_100a engine + C-ABI), the C++ drop-in layer and CLI when Eigen headers are
# reachable, and the test oracle.  nvcc cross-compiles without a GPU.
NVCC      ?= nvcc
CXX       ?= g++
ARCH      := -gencode arch=compute_100a,code=sm_100a
NVFLAGS   := $(ARCH) -O3 -std=c++17 -lineinfo -Xcompiler -fPIC -Xcompiler -Wall -Iinclude
CSRC      := pls_b200/csrc
CU        := $(CSRC)/api.cu $(CSRC)/dataset.cu $(CSRC)/pass.cu $(CSRC)/xty.cu $(CSRC)/scores.cu $(CSRC)/step.cu $(CSRC)/synth.cu $(CSRC)/gemm.cu $(CSRC)/allreduce.cu $(CSRC)/stats.cu \
             $(CSRC)/fit.cu $(wildcard $(CSRC)/predict.cu) $(wildcard $(CSRC)/cv.cu) $(CSRC)/cv_small.cu $(CSRC)/cv_gemm.cu $(CSRC)/noalloc.cu
OBJ       := $(patsubst $(CSRC)/%.cu,build/%.o,$(CU)) build/host_stats.o
HDR       := $(wildcard $(CSRC)/*.h $(CSRC)/*.cuh) include/plsb.h
LIB       := pls_b200/libplsb.so
EIGEN_INC ?= $(firstword $(wildcard /usr/include/eigen3 /root/reference/lib/eigen))

all: $(LIB) host oracle

$(LIB): $(OBJ)
	$(NVCC) $(ARCH) -shared -o $@ $(OBJ) -ldl

build/host_stats.o: $(CSRC)/host_stats.cpp include/plsb.h
	@mkdir -p build
	$(CXX) -std=c++17 -O2 -fPIC -Wall -c $< -o $@

build/%.o: $(CSRC)/%.cu $(HDR)
	@mkdir -p build
	$(NVCC) $(NVFLAGS) -c $< -o $@

HOSTSRC := pls_b200/host/model.cpp pls_b200/host/stats.cpp pls_b200/host/engine.cpp
ifneq ($(EIGEN_INC),)
# The C++ drop-in layer needs Eigen headers (the reference's own public dependency).  They are not in this
# repo: EIGEN_INC points at a system Eigen or at the reference checkout's submodule.  On a box without
# Eigen the prebuilt libpls.so / PLS / api_check (shipped with the snapshot) are used as they are.
HOSTFLAGS := -std=c++17 -O2 -fPIC -Wall -Iinclude -I$(EIGEN_INC)
host: pls_b200/libpls.so pls_b200/PLS tests/cpp/api_check tests/cpp/api_bench
pls_b200/libpls.so: $(HOSTSRC) pls_b200/host/engine.h include/PLS/pls.h include/plsb.h $(LIB)
	$(CXX) $(HOSTFLAGS) -shared -o $@ $(HOSTSRC) -Lpls_b200 -lplsb -lpthread -Wl,-rpath,'$$ORIGIN'
pls_b200/PLS: pls_b200/host/main.cpp pls_b200/libpls.so
	$(CXX) $(HOSTFLAGS) -o $@ pls_b200/host/main.cpp -Lpls_b200 -lpls -lplsb -lpthread -Wl,-rpath,'$$ORIGIN'
tests/cpp/api_check: tests/cpp/api_check.cpp pls_b200/libpls.so
	$(CXX) $(HOSTFLAGS) -o $@ tests/cpp/api_check.cpp -Lpls_b200 -lpls -lplsb -lpthread -Wl,-rpath,'$$ORIGIN/../../pls_b200'
tests/cpp/api_bench: tests/cpp/api_bench.cpp pls_b200/libpls.so
	$(CXX) $(HOSTFLAGS) -o $@ tests/cpp/api_bench.cpp -Lpls_b200 -lpls -lplsb -lpthread -Wl,-rpath,'$$ORIGIN/../../pls_b200'
else
host:
	@echo "host: no Eigen headers found (set EIGEN_INC); using prebuilt libpls.so / PLS if present"
endif

oracle:
	$(MAKE) -C oracle

clean:
	rm -rf build $(LIB) pls_b200/libpls.so pls_b200/PLS tests/cpp/api_check tests/cpp/api_bench
.PHONY: all host oracle clean
