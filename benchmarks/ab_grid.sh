#!/usr/bin/env bash
# ab_grid.sh <rounds> <which: a|b> <lib suffix>...: config 4 alone (benchmarks/grid_step.py), interleaved rounds
rounds=$1; which=$2; shift 2
mkdir -p gpurun_out
for r in $(seq 1 "$rounds"); do
  for n in "$@"; do
    lib=$PWD/cuda-raytracing_b200/lib/libb200rt$n.so
    [ -f "$lib" ] || continue
    B200RT_LIB=$lib timeout 600 python benchmarks/grid_step.py --steps 5 --which $which 2>/dev/null | tail -1 | sed "s/^{/{\"lib\": \"${n:-default}\", /" | tee -a gpurun_out/ab_grid.jsonl
  done
done
