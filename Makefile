# Top-level build: everything lands in-tree (git-ignored *.so, shipped by gpurun).
#
#   sift3d_b200/lib/libsift3d_cuda.so   hand-written sm_100a kernels + the CUDA C ABI
#   sift3d_b200/lib/libsift3D.so        host C: the reference's 27-symbol public API
#   sift3d_b200/lib/libsift3d_synth.so  synthetic workload generator (tests / bench)
#   oracle/...                          CPU oracle (test infrastructure), see oracle/Makefile

NVCC    ?= /usr/local/cuda/bin/nvcc
HOSTCC  := $(firstword $(wildcard /usr/bin/gcc) gcc)
HOSTCXX := $(firstword $(wildcard /usr/bin/g++) g++)
ARCH    := -gencode arch=compute_100a,code=sm_100a
# --fmad=false: the reference build never contracts a*b+c (x86-64 without -mfma); the
# kernels additionally use explicit __fmul_rn/__fadd_rn where bit parity is claimed.
NVFLAGS := $(ARCH) -O3 -lineinfo --fmad=false -std=c++17 -ccbin $(HOSTCXX) \
           -Xcompiler -fPIC,-fvisibility=hidden,-Wall,-Wno-unused-function -Xptxas -v
# descriptor.cu: bit parity is not claimed for descriptors (tolerance: cosine >= 0.999); the
# expressions that decide voxel membership / cell / face use explicit __f*_rn intrinsics and
# are unaffected, everything else may contract (later flag wins).
NVFLAGS_descriptor := --fmad=true
LIBDIR  := sift3d_b200/lib
CSRC    := sift3d_b200/csrc
CU_SRCS := $(CSRC)/sift3d_cuda.cu $(CSRC)/blur.cu $(CSRC)/extrema.cu $(CSRC)/orient.cu \
           $(CSRC)/descriptor.cu $(CSRC)/multi.cu
CU_OBJS := $(patsubst $(CSRC)/%.cu,build/%.o,$(CU_SRCS))
CU_HDRS := $(wildcard $(CSRC)/*.cuh) include/sift3d_cuda.h

all: product oracle

product: $(LIBDIR)/libsift3d_cuda.so $(LIBDIR)/libsift3D.so $(LIBDIR)/libsift3d_synth.so sift3d_b200/bin/kpSift3D

build/%.o: $(CSRC)/%.cu $(CU_HDRS)
	@mkdir -p build
	$(NVCC) $(NVFLAGS) $(NVFLAGS_$*) -c $< -o $@ 2> build/$*.ptxas.log || (cat build/$*.ptxas.log; false)

$(LIBDIR)/libsift3d_cuda.so: $(CU_OBJS)
	@mkdir -p $(LIBDIR)
	$(NVCC) $(ARCH) -shared -ccbin $(HOSTCXX) -o $@ $(CU_OBJS) -lpthread

$(LIBDIR)/libsift3D.so: sift3d_b200/host/sift3d_host.c sift3d_b200/host/nifti1.c sift3d_b200/host/nifti1.h include/sift3d.h include/sift3d_cuda.h $(LIBDIR)/libsift3d_cuda.so
	$(HOSTCC) -O2 -ffp-contract=off -fno-fast-math -fPIC -shared -fvisibility=hidden -Wall \
	    -o $@ sift3d_b200/host/sift3d_host.c sift3d_b200/host/nifti1.c -L$(LIBDIR) -lsift3d_cuda -lz -lm -lpthread \
	    -Wl,-rpath,'$$ORIGIN'

$(LIBDIR)/libsift3d_synth.so: sift3d_b200/synth.c
	@mkdir -p $(LIBDIR)
	$(HOSTCC) -O2 -ffp-contract=off -fno-fast-math -fopenmp -fPIC -shared -fvisibility=hidden \
	    -Wall -o $@ $< -lm

# The reference's command-line program (cli/kpSift3D.c), rewritten against the drop-in API.
sift3d_b200/bin/kpSift3D: cli/kpSift3D.c include/sift3d.h $(LIBDIR)/libsift3D.so
	@mkdir -p sift3d_b200/bin
	$(HOSTCC) -O2 -Wall -Iinclude -o $@ cli/kpSift3D.c -L$(LIBDIR) -lsift3D \
	    -Wl,-rpath,'$$ORIGIN/../lib'

oracle:
	$(MAKE) -C oracle all

clean:
	rm -rf build $(LIBDIR)/*.so sift3d_b200/bin
	$(MAKE) -C oracle clean

.PHONY: all product oracle clean
