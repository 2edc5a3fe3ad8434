# Builds the C-ABI library (sm_100a only) and the CPU oracle.  Used by __graft_entry__.build().
NVCC ?= /usr/local/cuda/bin/nvcc
PKG := aca-dotnet-workshop_b200
CSRC := $(PKG)/csrc
LIB := $(PKG)/libtasktable.so
NVFLAGS := -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -std=c++17 -Xcompiler -fPIC,-Wall \
           -Xptxas -v --expt-relaxed-constexpr
HDRS := $(wildcard $(CSRC)/*.cuh $(CSRC)/*.inc $(CSRC)/*.h include/*.h)

all: $(LIB) oracle/liboracle.so $(PKG)/host/tasks_manager_demo

$(LIB): $(CSRC)/tasktable.cu $(HDRS)
	$(NVCC) $(NVFLAGS) -shared -o $@ $(CSRC)/tasktable.cu 2> $(CSRC)/ptxas.log || (cat $(CSRC)/ptxas.log; false)

oracle/liboracle.so: oracle/oracle.cpp include/tt_synth.h include/tt_bmp_defined.h
	g++ -O2 -std=c++17 -fPIC -shared -pthread -o $@ oracle/oracle.cpp

$(PKG)/host/tasks_manager_demo: $(wildcard $(PKG)/host/*.cpp $(PKG)/host/*.hpp) include/tasktable.h
	@if [ -f $(PKG)/host/tasks_manager_demo.cpp ]; then \
	  g++ -O2 -std=c++17 -Iinclude -o $@ $(PKG)/host/tasks_manager_demo.cpp -ldl -pthread; \
	else true; fi

clean:
	rm -f $(LIB) oracle/liboracle.so $(CSRC)/ptxas.log $(PKG)/host/tasks_manager_demo

.PHONY: all clean
