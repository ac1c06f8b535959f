# Builds libg16b200.so (CUDA, sm_100a) and the `prover` CLI in-tree.  nvcc cross-compiles without a GPU.
PKG      := zk-proof-of-assets_b200
CSRC     := $(PKG)/csrc
NVCC     ?= nvcc
ARCH     := -gencode arch=compute_100a,code=sm_100a
NVFLAGS  := $(ARCH) -O3 -std=c++17 -lineinfo -Xcompiler -fPIC -Xcompiler -Wall -Iinclude
OBJS     := $(CSRC)/prover.o $(CSRC)/ops.o $(CSRC)/msm.o $(CSRC)/poseidon.o $(CSRC)/verify.o $(CSRC)/merkle.o $(CSRC)/setup.o $(CSRC)/zkey_new.o $(CSRC)/wtns_check.o
HDRS     := $(wildcard $(CSRC)/*.cuh) $(filter-out $(CSRC)/service% $(CSRC)/verify% $(CSRC)/binfile_ro%,$(wildcard $(CSRC)/*.hpp)) include/g16_b200.h

all: $(PKG)/libg16b200.so $(PKG)/libg16synth.so $(PKG)/prover

$(CSRC)/%.o: $(CSRC)/%.cu $(HDRS)
	$(NVCC) $(NVFLAGS) -c $< -o $@

$(CSRC)/zkey_new.o $(CSRC)/wtns_check.o: $(CSRC)/binfile_ro.hpp

$(CSRC)/verify.o: $(CSRC)/verify.cpp $(wildcard $(CSRC)/verify*.hpp) $(CSRC)/hostfield.hpp include/g16_b200.h
	g++ -O2 -std=c++17 -fPIC -Wall -Iinclude -c $< -o $@

$(CSRC)/merkle.o: $(CSRC)/merkle.cpp $(wildcard $(CSRC)/verify*.hpp) $(CSRC)/hostfield.hpp include/g16_b200.h
	g++ -O2 -std=c++17 -fPIC -Wall -Iinclude -c $< -o $@

$(CSRC)/setup.o: $(CSRC)/setup.cpp $(CSRC)/setup_hash.hpp $(wildcard $(CSRC)/verify*.hpp) $(CSRC)/hostfield.hpp include/g16_b200.h
	g++ -O2 -std=c++17 -fPIC -Wall -Iinclude -c $< -o $@

$(PKG)/libg16b200.so: $(OBJS)
	$(NVCC) $(ARCH) -shared -o $@ $(OBJS) -lcudart -lpthread

$(PKG)/libg16synth.so: $(CSRC)/synth.o
	$(NVCC) $(ARCH) -shared -o $@ $< -lcudart

$(PKG)/prover: $(CSRC)/prover_main.cpp $(wildcard $(CSRC)/service*.hpp) include/g16_b200.h $(PKG)/libg16b200.so
	g++ -O2 -std=c++17 -Wall -Iinclude $< -o $@ -L$(PKG) -lg16b200 -lpthread -Wl,-rpath,'$$ORIGIN'

clean:
	rm -f $(CSRC)/*.o $(PKG)/libg16b200.so $(PKG)/libg16synth.so $(PKG)/prover

.PHONY: all clean
