#!/bin/bash
# builds profiles/microbench/libmcb200_timing.so with the clock markers of the sweep / resident UDT kernels
set -e
cd /root/repo/montecarlo.jl_b200/csrc
mkdir -p /tmp/tb
ARCH="-gencode arch=compute_100a,code=sm_100a"
for f in sweep_delayed udt_resident segment_resident udt_blocked; do
  nvcc -O3 -std=c++17 -lineinfo $ARCH -Xcompiler -fPIC -DSWD_TIMING -DUDTR_TIMING -DMCB_EXPERIMENTS -c $f.cu -o /tmp/tb/$f.o
done
OBJS=""
for f in prof gemm gemm_tma udt udt_cluster rdivp_blocked sweep comm segment stab unequal measure dqmc_engine api_dqmc api_repulsive api_linalg ising; do OBJS="$OBJS $f.o"; done
nvcc -shared $ARCH -o /root/repo/profiles/microbench/libmcb200_timing.so $OBJS /tmp/tb/sweep_delayed.o /tmp/tb/udt_resident.o /tmp/tb/segment_resident.o /tmp/tb/udt_blocked.o -lcudart_static -lpthread -ldl -lrt
