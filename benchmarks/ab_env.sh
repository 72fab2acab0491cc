#!/usr/bin/env bash
# A/B of run-time switches: ab_env.sh VAR "v1 v2 ..."  -> bench.py step time per value (bunny 1080p 64 spp)
mkdir -p gpurun_out
var=$1; shift
for v in $1; do
  b=$(env $var=$v timeout 300 python bench.py --steps 8 --warmup 3 --no-cpu-baseline --no-e2e 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read()); print(round(d['ms_per_step'],2))")
  echo "$var=$v bench_ms=$b" | tee -a gpurun_out/ab_env.txt
done
