# Builds librtb200.so (the product: host scene API + sm_100a kernels behind the C ABI) and the
# CPU oracle (test infrastructure).  `python -c "import __graft_entry__ as g; g.build()"` runs this.
NVCC ?= nvcc
CSRC := raytracing_rust_b200/csrc
LIB ?= raytracing_rust_b200/librtb200.so
# experiment builds: make B=build/x LIB=build/x/librtb200.so EXTRA=-DRTB_FOO=1 lib
B ?= build
EXTRA ?=
NVEXTRA ?=
GENCODE := -gencode arch=compute_100a,code=sm_100a
NVCCFLAGS := -O3 -std=c++17 $(GENCODE) -lineinfo -Xcompiler -fPIC,-Wall,-Wno-unknown-pragmas -Iinclude -I$(CSRC) \
             -Xptxas -v --expt-relaxed-constexpr $(EXTRA) $(NVEXTRA)
HOST_SRCS := $(CSRC)/scene.cpp $(CSRC)/scenes_named.cpp $(CSRC)/flatten.cpp $(CSRC)/image_io.cpp
CU_SRCS := $(CSRC)/render.cu $(CSRC)/wavefront.cu $(CSRC)/peer.cu $(CSRC)/multi.cu $(CSRC)/calib.cu $(CSRC)/bvh_build.cu $(CSRC)/debug.cu
HDRS := $(wildcard include/*.h include/*.hpp $(CSRC)/*.h $(CSRC)/*.hpp $(CSRC)/*.cuh)
OBJS := $(patsubst $(CSRC)/%.cpp,$(B)/%.o,$(HOST_SRCS)) $(patsubst $(CSRC)/%.cu,$(B)/%.cu.o,$(CU_SRCS))

all: $(LIB) oracle
lib: $(LIB)

$(B)/%.o: $(CSRC)/%.cpp $(HDRS)
	@mkdir -p $(B)
	g++ -O2 -std=c++17 -fPIC -Wall -Wextra -Iinclude -I$(CSRC) $(EXTRA) -c $< -o $@

$(B)/%.cu.o: $(CSRC)/%.cu $(HDRS)
	@mkdir -p $(B)
	$(NVCC) $(NVCCFLAGS) -c $< -o $@ 2> $(B)/$*.ptxas.log || (cat $(B)/$*.ptxas.log; false)

$(LIB): $(OBJS)
	$(NVCC) -shared $(GENCODE) -o $@ $(OBJS) -cudart static

oracle:
	$(MAKE) -C oracle

clean:
	rm -rf build $(LIB)
	$(MAKE) -C oracle clean

.PHONY: all lib oracle clean
