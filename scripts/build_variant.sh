#!/bin/bash
# Development aid: build/variants/libfeasm_<name>.so with extra -D flags (ablation / A-B builds; FEASM_LIB selects one)
# usage: scripts/build_variant.sh name -DFLAG1 -DFLAG2 ...
set -e
name=$1; shift
cd "$(dirname "$0")/../libmesh_b200/csrc"
out=../../build/variants/$name
mkdir -p $out
for f in feasm_api.cu feasm_element.cu feasm_scatter.cu feasm_solve.cu feasm_side.cu feasm_pattern.cu feasm_ingest.cu feasm_peaks.cu feasm_noalloc.cu feasm_comm.cu feasm_general.cu feasm_tables.cpp feasm_host.cpp feasm_plan.cpp; do
  if [ $f = feasm_element.cu ] || [ $f = feasm_scatter.cu ] || [ ! -f $out/$f.o ] || [ $f -nt $out/$f.o ] || [ feasm_internal.h -nt $out/$f.o ]; then
    nvcc -O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -lineinfo -Xcompiler -fPIC -I../../include -I. -x cu "$@" -c $f -o $out/$f.o &
  fi
done
wait
nvcc -gencode arch=compute_100a,code=sm_100a -shared -o ../../build/variants/libfeasm_$name.so $out/*.o -lcudart -ldl
echo built build/variants/libfeasm_$name.so
