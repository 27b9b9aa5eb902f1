#!/bin/bash
# usage: tools/ab_build.sh <name> <src.cu> <extra nvcc flags...>: builds build_ab/lib_<name>.so with one source recompiled
# (select it at run time with HYDRA_GNN_B200_SO=/root/repo/hydra_gnn_b200/ab/lib_<name>.so)
name=$1; src=$2; shift 2
cd /root/repo/hydra_gnn_b200/csrc || exit 1
mkdir -p /root/repo/hydra_gnn_b200/ab
base=$(basename $src .cu)
nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -Xcompiler -fPIC,-fvisibility=hidden -I/root/repo/include --expt-relaxed-constexpr "$@" -c -o /root/repo/hydra_gnn_b200/ab/${base}_$name.o $src || exit 1
objs=""
for o in $(sed -n 's/^SRCS := //p' Makefile | sed 's/\.cu//g'); do
  if [ $o = $base ]; then objs="$objs /root/repo/hydra_gnn_b200/ab/${base}_$name.o"; else objs="$objs build/$o.o"; fi
done
nvcc -gencode arch=compute_100a,code=sm_100a -shared -o /root/repo/hydra_gnn_b200/ab/lib_$name.so $objs -cudart static -lcuda && rm -f /root/repo/hydra_gnn_b200/ab/${base}_$name.o
