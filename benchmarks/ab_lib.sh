#!/usr/bin/env bash
# A/B of build variants on the bench step only: ab_lib.sh <lib suffix>...   ("" = default build); also checks the frame is bit-identical
mkdir -p gpurun_out
for n in "$@"; do
  lib=$PWD/cuda-raytracing_b200/lib/libb200rt$n.so
  [ -f "$lib" ] || continue
  b=$(B200RT_LIB=$lib timeout 300 python bench.py --steps 8 --warmup 3 --no-cpu-baseline --no-e2e 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read()); print(round(d['ms_per_step'],2))")
  h=$(B200RT_LIB=$lib timeout 300 python benchmarks/frame_hash.py 2>/dev/null | tail -1)
  echo "variant=${n:-default} bench_ms=$b frame_hash=$h" | tee -a gpurun_out/ab_lib.txt
done
