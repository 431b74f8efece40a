# Product build: mvolps_b200/lib/libglpk_b200.so (C++ host + sm_100a kernels)
# and the CLI mvolps_b200/bin/mvolps_b200.  `make oracle` builds the test-only
# CPU oracle (see oracle/Makefile).
ROOT := $(abspath $(dir $(lastword $(MAKEFILE_LIST))))
CSRC := $(ROOT)/mvolps_b200/csrc
OBJ  := $(ROOT)/build
LIB  := $(ROOT)/mvolps_b200/lib/libglpk_b200.so
BIN  := $(ROOT)/mvolps_b200/bin/mvolps_b200
NVCC := nvcc
ARCH := -gencode arch=compute_100a,code=sm_100a
NVFLAGS := $(ARCH) -lineinfo -O3 -std=c++17 -fmad=false -Xcompiler -fPIC -I$(ROOT)/include $(if $(PROF),-DMVB_PROF) $(if $(PSNT),-DMVB_PS_NT=$(PSNT))
CXXFLAGS := -std=c++17 -O2 -mfma -ffp-contract=off -fPIC -Wall -I$(ROOT)/include
HOST := glpk_shim lp_reader bnb_driver

all: $(LIB) $(BIN)

$(OBJ)/engine_cuda.o: $(CSRC)/engine_cuda.cu $(wildcard $(CSRC)/*.cuh) $(ROOT)/include/mvolps_engine.h $(ROOT)/include/mvolps_spec.h $(ROOT)/include/mvolps_dist.h
	@mkdir -p $(OBJ)
	$(NVCC) $(NVFLAGS) -c $< -o $@

# (MVB_WITH_SHARDS: the product's glp_simplex can route a large LP to the column-sharded engine; the
# oracle build of the same host sources has no multi-GPU layer)
$(OBJ)/%.o: $(CSRC)/%.cpp $(CSRC)/lp_model.hpp $(CSRC)/cow_vec.hpp $(CSRC)/par_for.hpp $(ROOT)/include/glpk.h $(ROOT)/include/mvolps_engine.h $(ROOT)/include/mvolps_bnb.h $(ROOT)/include/mvolps_dist.h
	@mkdir -p $(OBJ)
	g++ $(CXXFLAGS) -DMVB_WITH_SHARDS -c $< -o $@

$(LIB): $(OBJ)/engine_cuda.o $(addprefix $(OBJ)/,$(addsuffix .o,$(HOST)))
	@mkdir -p $(dir $@)
	$(NVCC) $(ARCH) -shared -o $@ $^ -Xlinker -Bsymbolic -ldl

$(BIN): $(OBJ)/mvolps_main.o $(LIB)
	@mkdir -p $(dir $@)
	g++ -o $@ $(OBJ)/mvolps_main.o -L$(dir $(LIB)) -lglpk_b200 -Wl,-rpath,'$$ORIGIN/../lib'

oracle:
	$(MAKE) -C $(ROOT)/oracle all
	@if [ -d /root/reference ]; then $(MAKE) -C $(ROOT)/oracle ref; fi

clean:
	rm -rf $(OBJ) $(ROOT)/mvolps_b200/lib $(ROOT)/mvolps_b200/bin

.PHONY: all oracle clean
