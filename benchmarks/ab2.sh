#!/usr/bin/env bash
# A/B of trace_wide structure switches (RT_TW_FOLD / RT_TW_MODE): parity subset + incoherent traversal + bench step per lib variant
# usage: ab2.sh "<minblocks list>" <lib suffix>...   ("" = the default library)
mkdir -p gpurun_out
mbs=${1:-5}; shift
for n in "" "$@"; do
  lib=$PWD/cuda-raytracing_b200/lib/libb200rt$n.so
  [ -f "$lib" ] || continue
  par=$(B200RT_LIB=$lib timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "primary or incoherent or fresh or 1080p or edge or path_trace_frames" 2>&1 | tail -1)
  inc=$(B200RT_LIB=$lib timeout 300 python benchmarks/incoherent.py 8 2>/dev/null | tail -1 | python -c "import json,sys; d=json.loads(sys.stdin.read()); print(round(d['ms_v0'],3))")
  for mb in $mbs; do
    b=$(B200RT_LIB=$lib B200RT_MINBLOCKS=$mb timeout 300 python bench.py --steps 8 --warmup 3 --no-cpu-baseline --no-e2e 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read()); print(round(d['ms_per_step'],2))")
    echo "variant=${n:-default} minblocks=$mb parity=[$par] incoherent_ms=$inc bench_ms_per_step=$b" | tee -a gpurun_out/ab2.txt
  done
done
