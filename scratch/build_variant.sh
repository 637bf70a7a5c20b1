#!/bin/bash
# build a variant of libdgibbs.so with extra -D flags:  scratch/build_variant.sh NAME -DDGIBBS_TKC=32 -DDGIBBS_TNST=2
# -> scratch/variants/libdgibbs_NAME.so  (select with DGIBBS_LIB_PATH); RECOMPILE="a b" names the sources the flags
# touch (default: sampler_dense), the other objects are taken from the in-tree build
set -e
cd "$(dirname "$0")/.."
name=$1; shift
mkdir -p scratch/variants/obj_$name
objs=""
for src in dgibbs_api setup_kernels sampler_dense sampler_diag sampler_cross sampler_freq project_fft cut_kernels finish_kernels delay_fft; do
  o=scratch/variants/obj_$name/$src.o
  if [[ " ${RECOMPILE:-sampler_dense} " == *" $src "* ]] || [ ! -f draco_b200/build/$src.o ]; then
    nvcc -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -std=c++17 -Xcompiler -fPIC "$@" -c draco_b200/csrc/$src.cu -o $o 2>&1 | grep -v "nvcc warning" || true
  else
    cp draco_b200/build/$src.o $o
  fi
  objs="$objs $o"
done
nvcc -shared -o scratch/variants/libdgibbs_$name.so $objs 2>&1 | grep -v "nvcc warning" || true
ls -la scratch/variants/libdgibbs_$name.so
