# Builds the B200 factorization library and the drop-in `matFact` program.
#
#   make            -> recommender_system_b200/libmatfact_b200.so, bin/matFact
#   make oracle     -> oracle/_build (CPU checker, test infrastructure) and, when
#                      /root/reference is present, oracle/_ref (unmodified reference)
#
# sm_100a only.  -fmad=false: the reference never fuses multiply and add
# (x86-64, matFact.c:45-51); the kernels use __dmul_rn/__dadd_rn explicitly and
# this flag is the second line of defence.
NVCC     ?= /usr/local/cuda/bin/nvcc
CC       = gcc
ARCH     = -gencode arch=compute_100a,code=sm_100a
NVFLAGS  = $(ARCH) -O3 -lineinfo -fmad=false -std=c++17 -Xcompiler -fPIC,-Wall,-Wextra -Iinclude
CFLAGS   = -O2 -ffp-contract=off -Wall -Wextra -fPIC -Iinclude
PKG      = recommender_system_b200
LIB      = $(PKG)/libmatfact_b200.so
CSRC     = $(PKG)/csrc
HDRS     = include/matfact_b200.h $(PKG)/host/mf_prep.h $(PKG)/host/mf_ingest.h $(CSRC)/mf_internal.h $(CSRC)/mf_types.h $(CSRC)/mf_loop.cuh $(CSRC)/mf_argmax.cuh
OBJS     = build/mf_capi.o build/mf_plan.o build/mf_iter.o build/mf_prep.o build/mf_ingest.o

all: $(LIB) bin/matFact

build/%.o: $(CSRC)/%.cu $(HDRS)
	@mkdir -p build
	$(NVCC) $(NVFLAGS) -c $< -o $@

build/mf_prep.o: $(PKG)/host/mf_prep.c $(PKG)/host/mf_prep.h include/matfact_b200.h
	@mkdir -p build
	$(CC) $(CFLAGS) -c $< -o $@

build/mf_ingest.o: $(PKG)/host/mf_ingest.c $(PKG)/host/mf_ingest.h include/matfact_b200.h
	@mkdir -p build
	$(CC) $(CFLAGS) -O3 -c $< -o $@

$(LIB): $(OBJS)
	$(NVCC) $(ARCH) -shared -o $@ $(OBJS) -ldl

# The host program stays plain C (north star) and only sees the C ABI.
bin/matFact: $(PKG)/host/matFact.c include/matfact_b200.h $(LIB)
	@mkdir -p bin
	$(CC) $(CFLAGS) -o $@ $(PKG)/host/matFact.c -L$(PKG) -lmatfact_b200 -Wl,-rpath,'$$ORIGIN/../$(PKG)'

oracle:
	$(MAKE) -C oracle all ref

ptxas-info:
	$(NVCC) $(NVFLAGS) -Xptxas -v -c $(CSRC)/mf_iter.cu -o /dev/null

clean:
	rm -rf build bin $(LIB)

.PHONY: all oracle clean ptxas-info
