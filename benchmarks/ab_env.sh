#!/usr/bin/env bash
# Interleaved A/B of builder / runtime switches read from the environment, on the GPU box:
#   ab_env.sh <rounds> "<extra ab_step args>" "NAME=VALUE [NAME=VALUE ...]" "..."      ("-" = no switches)
# e.g. ab_env.sh 2 "--grid" "-" "B200RT_TOPOLOGY=ref". One fresh process per measurement, rounds interleave the settings.
rounds=$1; extra=$2; shift 2
mkdir -p gpurun_out
for r in $(seq 1 "$rounds"); do
  for setting in "$@"; do
    envs=$setting; [ "$setting" = "-" ] && envs=""
    env $envs timeout 900 python benchmarks/ab_step.py $extra 2>/dev/null | tail -1 | sed "s/^{/{\"round\": $r, \"env\": \"$setting\", /" | tee -a gpurun_out/ab_env.jsonl
  done
done
