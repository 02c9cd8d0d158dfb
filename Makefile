# Builds libvst_b200.so (sm_100a only) in-tree.  `python -c "import __graft_entry__ as g; g.build()"`
# runs the same commands.
NVCC      ?= nvcc
ARCH      := -gencode arch=compute_100a,code=sm_100a
NVFLAGS   := $(ARCH) -O3 -std=c++17 -lineinfo -Xcompiler -fPIC -Xptxas -v --expt-relaxed-constexpr
CSRC      := video_style_transfer_b200/csrc
SRCS      := $(wildcard $(CSRC)/*.cu)
OBJS      := $(SRCS:.cu=.o)
LIB       := video_style_transfer_b200/libvst_b200.so

all: $(LIB)

$(CSRC)/%.o: $(CSRC)/%.cu $(CSRC)/common.cuh include/vst_b200.h
	$(NVCC) $(NVFLAGS) -c $< -o $@

$(LIB): $(OBJS)
	$(NVCC) $(ARCH) -shared -o $@ $(OBJS)

clean:
	rm -f $(OBJS) $(LIB)

# microbenchmarks behind the notes in profiles/ (MMA issue rate, MUFU rate); not part of the library
ubench: scripts/ubench/mma_rate.bin scripts/ubench/mufu_rate.bin
scripts/ubench/%.bin: scripts/ubench/%.cu
	$(NVCC) $(ARCH) -O3 -std=c++17 -o $@ $<
