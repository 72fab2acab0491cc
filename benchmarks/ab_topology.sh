#!/usr/bin/env bash
# Interleaved A/B of the wide tree's topology on the GPU box: ab_topology.sh <rounds> <extra ab_step args or "">
# B200RT_TOPOLOGY=ref = the reference's midpoint-split tree re-laid out as it is; unset = the SAH tree over the same leaves.
rounds=$1; extra=$2
mkdir -p gpurun_out
for r in $(seq 1 "$rounds"); do
  for t in sah ref; do
    B200RT_TOPOLOGY=$t timeout 900 python benchmarks/ab_step.py $extra 2>/dev/null | tail -1 | sed "s/^{/{\"round\": $r, \"topology\": \"$t\", /" | tee -a gpurun_out/ab_topology.jsonl
  done
done
