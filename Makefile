# Builds libfourier_b200.so (sm_100a CUDA kernels + C ABI), the C oracle and the
# host-arithmetic test harness.  `python __graft_entry__.py build` runs this.
NVCC      ?= nvcc
CXX       ?= g++
CC        := /usr/bin/gcc
ARCH      := -gencode arch=compute_100a,code=sm_100a
NVFLAGS   := $(ARCH) -O3 -std=c++17 -lineinfo -Xcompiler -fPIC -Xcompiler -O2 --expt-relaxed-constexpr -Xptxas -v $(EXTRA_NVFLAGS)
CSRC      := fourier_b200/csrc
BUILD     := build
CU        := ntt poly msm g1ops srs api engine verify evalform
OBJS      := $(addprefix $(BUILD)/,$(addsuffix .o,$(CU))) $(BUILD)/g2host.o
HDRS      := $(wildcard $(CSRC)/*.cuh) include/fourier_b200.h
LIB       := fourier_b200/libfourier_b200.so

all: $(LIB) oracle tests/_build/libhost_arith.so

$(BUILD)/%.o: $(CSRC)/%.cu $(HDRS)
	@mkdir -p $(BUILD)
	$(NVCC) $(NVFLAGS) -c $< -o $@ 2> $(BUILD)/$*.ptxas.log || (cat $(BUILD)/$*.ptxas.log; false)

$(BUILD)/g2host.o: $(CSRC)/g2host.cpp $(HDRS)
	@mkdir -p $(BUILD)
	$(CXX) -O2 -std=c++17 -fPIC -x c++ -I$(CSRC) -c $< -o $@

$(LIB): $(OBJS)
	$(NVCC) $(ARCH) -shared -o $@ $(OBJS) -cudart static

oracle: oracle/c/liboracle.so

oracle/c/liboracle.so: oracle/c/oracle.c
	$(CC) -O3 -mbmi2 -madx -fopenmp -shared -fPIC -o $@ $<

tests/_build/libhost_arith.so: tests/host_arith.cpp $(HDRS)
	@mkdir -p tests/_build
	$(CXX) -O2 -std=c++17 -x c++ -shared -fPIC -o $@ $<

clean:
	rm -rf $(BUILD) $(LIB) oracle/c/liboracle.so tests/_build

.PHONY: all oracle clean
