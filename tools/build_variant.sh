#!/bin/sh
# tools/build_variant.sh NAME "-DMF_...=.. ..." : libmatfact_b200_NAME.so with extra defines for mf_iter.cu
# (development aid: MFGPU_LIB=<path> makes api.py load it instead of the in-tree library)
set -e
N=$1; D=$2
mkdir -p build/v_$N
/usr/local/cuda/bin/nvcc -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -fmad=false -std=c++17 -Xcompiler -fPIC -Iinclude $D -c recommender_system_b200/csrc/mf_iter.cu -o build/v_$N/mf_iter.o
/usr/local/cuda/bin/nvcc -gencode arch=compute_100a,code=sm_100a -shared -o recommender_system_b200/libmatfact_b200_$N.so build/mf_capi.o build/mf_plan.o build/v_$N/mf_iter.o build/mf_prep.o build/mf_ingest.o -ldl
