#!/usr/bin/env bash
# Interleaved A/B rounds on the GPU box: ab_round.sh <rounds> <extra ab_step args or ""> <lib suffix>...   ("" = default build)
# Every variant is measured once per round (a fresh process each), rounds interleave the variants so drift hits all alike.
rounds=$1; extra=$2; shift 2
mkdir -p gpurun_out
for r in $(seq 1 "$rounds"); do
  for n in "$@"; do
    lib=$PWD/cuda-raytracing_b200/lib/libb200rt$n.so
    [ -f "$lib" ] || continue
    B200RT_LIB=$lib timeout 600 python benchmarks/ab_step.py $extra 2>/dev/null | tail -1 | sed "s/^{/{\"round\": $r, /" | tee -a gpurun_out/ab_round.jsonl
  done
done
