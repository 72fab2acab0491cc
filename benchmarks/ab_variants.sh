#!/usr/bin/env bash
# ab_variants.sh <rounds> <extra ab_step args or ""> <variant>...: kernel variants of ONE build, interleaved rounds, a fresh process each
rounds=$1; extra=$2; shift 2
mkdir -p gpurun_out
for r in $(seq 1 "$rounds"); do
  for v in "$@"; do
    timeout 600 python benchmarks/ab_step.py --variant $v $extra 2>/dev/null | tail -1 | sed "s/^{/{\"round\": $r, /" | tee -a gpurun_out/ab_variants.jsonl
  done
done
