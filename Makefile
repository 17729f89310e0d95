# Builds the product library (sm_100a only) and, through oracle/Makefile, the test oracle.
PKG := ray-tracer-challenge_b200
CSRC := $(PKG)/csrc
NVCC ?= nvcc
# -fmad=false: the reference (Zig, no fast-math) never fuses a*b+c; keeping IEEE mul+add makes t values,
# tie-breaks and pattern cell decisions bit-identical to it (DESIGN.md "exact arithmetic").
FMAD ?= false
EXTRA ?=
NVFLAGS ?= $(EXTRA) -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -fmad=$(FMAD) -Xcompiler -fPIC,-Wall
INST := $(foreach m,0 1,$(foreach f,0 2 3,$(CSRC)/render_inst_$(m)_$(f).cu))
SRC := $(INST) $(CSRC)/render_mesh.cu $(CSRC)/render_analytic.cu $(CSRC)/tile_order.cu $(CSRC)/bvh_gpu.cu $(CSRC)/rtc_api.cu $(CSRC)/scene_build.cpp $(CSRC)/bvh.cpp \
       $(CSRC)/host/scene.cpp $(CSRC)/host/scene_parser.cpp $(CSRC)/host/obj.cpp $(CSRC)/host/images.cpp \
       $(CSRC)/host/flatten.cpp $(CSRC)/rtc_renderer.cpp $(CSRC)/host/image_write.cpp
HDR := $(wildcard $(CSRC)/*.h $(CSRC)/*.hpp $(CSRC)/*.cuh $(CSRC)/host/*.hpp include/*.h)

OUT ?= $(PKG)/librtc_b200.so

all: $(OUT) oracle tools/peaks

# one object per source so that `make -j` compiles the kernel instantiations in parallel
OBJDIR ?= build/$(notdir $(basename $(OUT)))
OBJ := $(patsubst $(CSRC)/%,$(OBJDIR)/%.o,$(SRC))
$(OBJDIR)/%.o: $(CSRC)/% $(HDR) Makefile
	@mkdir -p $(dir $@)
	$(NVCC) $(NVFLAGS) -c -o $@ $<
$(OUT): $(OBJ)
	$(NVCC) $(NVFLAGS) -shared -o $@ $(OBJ) -lz -lpthread

oracle:
	$(MAKE) -C oracle all

# measured FP64 / FP32 FMA peak and L2 copy bandwidth (denominators of the ALU / L2 rooflines)
tools/peaks: tools/peaks.cu
	$(NVCC) -O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -lineinfo -o $@ $<

clean:
	rm -f $(PKG)/librtc_b200.so
	$(MAKE) -C oracle clean

.PHONY: all oracle clean
