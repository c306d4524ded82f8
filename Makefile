# Builds the product library (hand-written sm_100a kernels + C ABI) in-tree, and the test oracle.
#   make            -> upmem--openfhe_b200/lib/libfhe_b200.so  and  oracle/libfhe_oracle.so
#   make ref        -> oracle/_ref/libopenfhe_ref.so (needs /root/reference; this container only)
NVCC    ?= /usr/local/cuda/bin/nvcc
PKG     := upmem--openfhe_b200
CSRC    := $(PKG)/csrc
LIBDIR  := $(PKG)/lib
NVFLAGS := -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 --expt-relaxed-constexpr -Xcompiler -fPIC -Iinclude -I$(CSRC)
CU      := $(CSRC)/elementwise.cu $(CSRC)/ntt.cu $(CSRC)/ntt_ws.cu $(CSRC)/modops.cu $(CSRC)/automorph.cu $(CSRC)/basis.cu $(CSRC)/keyswitch.cu $(CSRC)/decrypt.cu $(CSRC)/behz.cu $(CSRC)/runtime.cu
OBJ     := $(patsubst $(CSRC)/%.cu,$(LIBDIR)/obj/%.o,$(CU))
HDR     := $(wildcard $(CSRC)/*.h $(CSRC)/*.cuh include/*.h include/pim/*.h)

.PHONY: all lib oracle ref host tools clean
all: lib oracle host tools
lib: $(LIBDIR)/libfhe_b200.so
$(LIBDIR)/obj/%.o: $(CSRC)/%.cu $(HDR)
	@mkdir -p $(dir $@)
	$(NVCC) $(NVFLAGS) -c $< -o $@
$(LIBDIR)/libfhe_b200.so: $(OBJ)
	$(NVCC) -gencode arch=compute_100a,code=sm_100a -shared -o $@ $^ -ldl
oracle:
	$(MAKE) -C oracle oracle
ref:
	$(MAKE) -C oracle ref
# C++ parity test of the include/pim surface (links the product .so and, as its checker, the oracle)
host: lib oracle tests/cpp/bin/pim_surface_test tests/cpp/bin/pim_multi_gpu_test
tests/cpp/bin/pim_multi_gpu_test: tests/cpp/pim_multi_gpu_test.cpp $(HDR)
	@mkdir -p $(dir $@)
	g++ -std=c++17 -O2 -Wall -Iinclude -Ioracle $< -o $@ -L$(LIBDIR) -lfhe_b200 -Loracle -lfhe_oracle \
	    -Wl,-rpath,'$$ORIGIN/../../../$(LIBDIR)' -Wl,-rpath,'$$ORIGIN/../../../oracle'
tests/cpp/bin/pim_surface_test: tests/cpp/pim_surface_test.cpp $(HDR)
	@mkdir -p $(dir $@)
	g++ -std=c++17 -O2 -Wall -Iinclude -Ioracle $< -o $@ -L$(LIBDIR) -lfhe_b200 -Loracle -lfhe_oracle \
	    -Wl,-rpath,'$$ORIGIN/../../../$(LIBDIR)' -Wl,-rpath,'$$ORIGIN/../../../oracle'
# measurement tools that run on the GPU box (built here, travel with the snapshot)
tools: tools/bin/int_probe
tools/bin/int_probe: tools/int_probe.cu
	@mkdir -p $(dir $@)
	$(NVCC) -gencode arch=compute_100a,code=sm_100a -O3 -o $@ $<
clean:
	rm -rf $(LIBDIR) tools/bin && $(MAKE) -C oracle clean
