# Builds libpfbwt_b200.so (sm_100a only) and the drop-in binaries pfbwt-f, pfbwt-f64, merge_pfp.
NVCC   ?= nvcc
CXX    ?= g++
ARCH   := -gencode arch=compute_100a,code=sm_100a
NVFLAGS := -std=c++17 -O3 $(ARCH) -lineinfo --extended-lambda -Xcompiler -fPIC,-Wno-deprecated-declarations \
           -Wno-deprecated-declarations -I pfbwt-f_b200/csrc -I include
CSRC   := pfbwt-f_b200/csrc
LIB    := pfbwt-f_b200/libpfbwt_b200.so
BIN    := bin

all: $(LIB) $(BIN)/pfbwt-f $(BIN)/pfbwt-f64 $(BIN)/merge_pfp $(BIN)/mps_to_ma $(BIN)/shim_demo $(BIN)/shim_demo64 emu

$(LIB): $(CSRC)/capi.cu $(CSRC)/engine.hpp $(CSRC)/pipeline.cuh $(CSRC)/build.cuh $(CSRC)/dictkeys.cuh $(CSRC)/emit.cuh $(CSRC)/radix.cuh $(CSRC)/prim.cuh $(CSRC)/host_io.hpp $(CSRC)/comm.hpp $(CSRC)/shard.inl $(CSRC)/ingest.cuh $(CSRC)/devio.cuh $(CSRC)/parsekernels.cuh $(CSRC)/scan.cuh $(CSRC)/markers.cuh include/pfbwt_b200.h
	$(NVCC) $(NVFLAGS) -shared -o $@ $(CSRC)/capi.cu -lz -lcudart -ldl -lpthread

$(BIN)/pfbwt-f: pfbwt-f_b200/src/pfbwt-f.cpp $(LIB)
	mkdir -p $(BIN)
	$(CXX) -std=c++17 -O2 -I include -o $@ $< -Lpfbwt-f_b200 -lpfbwt_b200 -Wl,-rpath,'$$ORIGIN/../pfbwt-f_b200'
$(BIN)/pfbwt-f64: pfbwt-f_b200/src/pfbwt-f.cpp $(LIB)
	mkdir -p $(BIN)
	$(CXX) -std=c++17 -O2 -DM64 -I include -o $@ $< -Lpfbwt-f_b200 -lpfbwt_b200 -Wl,-rpath,'$$ORIGIN/../pfbwt-f_b200'
$(BIN)/merge_pfp: pfbwt-f_b200/src/merge_pfp.cpp $(LIB)
	mkdir -p $(BIN)
	$(CXX) -std=c++17 -O2 -I include -o $@ $< -Lpfbwt-f_b200 -lpfbwt_b200 -Wl,-rpath,'$$ORIGIN/../pfbwt-f_b200'

$(BIN)/mps_to_ma: pfbwt-f_b200/src/mps_to_ma.cpp $(LIB)
	mkdir -p $(BIN)
	$(CXX) -std=c++17 -O2 -I include -o $@ $< -Lpfbwt-f_b200 -lpfbwt_b200 -Wl,-rpath,'$$ORIGIN/../pfbwt-f_b200'

# a user-style driver compiled against the source-level shim of the reference's C++ API (tests/test_shim.py)
$(BIN)/shim_demo: tests/shim/shim_main.cpp include/pfbwtf_shim.hpp $(LIB)
	mkdir -p $(BIN)
	$(CXX) -std=c++17 -O2 -Wall -I include -o $@ $< -Lpfbwt-f_b200 -lpfbwt_b200 -Wl,-rpath,'$$ORIGIN/../pfbwt-f_b200'
$(BIN)/shim_demo64: tests/shim/shim_main.cpp include/pfbwtf_shim.hpp $(LIB)
	mkdir -p $(BIN)
	$(CXX) -std=c++17 -O2 -Wall -DM64=1 -I include -o $@ $< -Lpfbwt-f_b200 -lpfbwt_b200 -Wl,-rpath,'$$ORIGIN/../pfbwt-f_b200'

# developer logic check (tests only; see tests/emu/emu_main.cpp)
emu: tests/emu/pfb_emu tests/emu/libpfb_emu.so
tests/emu/libpfb_emu.so: tests/emu/pfb_emu
	$(CXX) -std=c++17 -O2 -Wall -Wno-unused -shared -fPIC -I $(CSRC) -o $@ tests/emu/emu_main.cpp -lz -pthread
tests/emu/pfb_emu: tests/emu/emu_main.cpp $(CSRC)/engine.hpp $(CSRC)/pipeline.cuh $(CSRC)/build.cuh $(CSRC)/dictkeys.cuh $(CSRC)/emit.cuh $(CSRC)/radix.cuh $(CSRC)/prim.cuh $(CSRC)/host_io.hpp $(CSRC)/comm.hpp $(CSRC)/shard.inl $(CSRC)/ingest.cuh $(CSRC)/devio.cuh $(CSRC)/parsekernels.cuh $(CSRC)/scan.cuh $(CSRC)/markers.cuh
	$(CXX) -std=c++17 -O2 -Wall -Wno-unused -I $(CSRC) -o $@ $< -lz -pthread

oracle:
	$(MAKE) -C oracle all

clean:
	rm -rf $(LIB) $(BIN) tests/emu/pfb_emu tests/emu/libpfb_emu.so

.PHONY: all emu oracle clean
