#!/usr/bin/env bash
# per lib variant and CTAs/SM: bench.py step (bunny, L2-resident) and the 13.9 M-triangle grid (config 4b, not L2-resident)
# usage: ab4.sh "<minblocks list>" <lib suffix>...
mkdir -p gpurun_out
mbs=${1:-5}; shift
for n in "$@"; do
  lib=$PWD/cuda-raytracing_b200/lib/libb200rt$n.so
  [ -f "$lib" ] || continue
  for mb in $mbs; do
    b=$(B200RT_MINBLOCKS=$mb B200RT_LIB=$lib timeout 300 python bench.py --steps 8 --warmup 3 --no-cpu-baseline --no-e2e 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read()); print(round(d['ms_per_step'],2))")
    g=$(B200RT_MINBLOCKS=$mb B200RT_LIB=$lib timeout 600 python benchmarks/scaled_meshes.py b 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read()); print(round(d['pt_ms'],2), round(d['primary_ms'],3), d['primary_ids_equal_reference_order'])")
    echo "variant=${n:-default} minblocks=$mb bench_ms=$b grid4b_pt_ms,primary_ms,ids_ok=$g" | tee -a gpurun_out/ab4.txt
  done
done
