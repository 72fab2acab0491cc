#!/usr/bin/env bash
# A/B of trace_wide tuning switches: for every lib variant, incoherent-ray traversal time and the bench step time
for n in "" _fma _cache _tos _all "$@"; do
  lib=cuda-raytracing_b200/lib/libb200rt$n.so
  [ -f "$lib" ] || continue
  inc=$(B200RT_LIB=$PWD/$lib python benchmarks/incoherent.py 8 2>/dev/null | tail -1 | python -c "import json,sys; d=json.loads(sys.stdin.read()); print(round(d['ms_v0'],3))")
  for mb in 4 5; do
    b=$(B200RT_LIB=$PWD/$lib B200RT_MINBLOCKS=$mb python bench.py --steps 5 --warmup 3 --no-cpu-baseline --no-e2e 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read()); print(round(d['ms_per_step'],2))")
    echo "variant=${n:-base} minblocks=$mb incoherent_ms=$inc bench_ms_per_step=$b"
  done
done
