#!/usr/bin/env bash
# On the GPU box, one step per gpurun call (at most one ncu run per call, and only after the same command exited 0 without ncu):
#   bash benchmarks/gpu_bench_and_profile.sh <tag> bench   bench line + reference arm                     -> gpurun_out/<tag>_bench*.json
#   bash benchmarks/gpu_bench_and_profile.sh <tag> list    ncu launch list of the bench command           -> gpurun_out/<tag>_launches.csv
#   bash benchmarks/gpu_bench_and_profile.sh <tag> full    one ncu --set full capture of the megakernel   -> gpurun_out/<tag>_megakernel.ncu-rep
tag=$1; mode=${2:-bench}
mkdir -p gpurun_out
P="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-e2e --no-strong --no-configs"
case $mode in
bench)
  python bench.py > gpurun_out/${tag}_bench.json 2> gpurun_out/${tag}_bench.err || { tail -30 gpurun_out/${tag}_bench.err; exit 1; }
  python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/${tag}_bench_reference_arm.json 2>/dev/null
  cat gpurun_out/${tag}_bench.json | python -c "import json,sys; d=json.loads(sys.stdin.read()); print({k: d[k] for k in ('value','ms_per_step','e2e','clocks')}); print(d['roofline']); print(d.get('strong')); print(d.get('configs'))"
  ;;
list)
  $P > gpurun_out/${tag}_plain.log 2>&1 &&
  ncu --metrics gpu__time_duration.sum --clock-control none -c 40 --csv --log-file gpurun_out/${tag}_launches.csv $P > gpurun_out/${tag}_ncu_list.log 2>&1
  tail -3 gpurun_out/${tag}_launches.csv
  ;;
full)
  $P --no-flush > gpurun_out/${tag}_plain2.log 2>&1 &&
  ncu --set full --clock-control none --import-source on --kernel-name-base mangled -k regex:pathtrace_megakernelILi0ELi5 -s 3 -c 1 \
      -o gpurun_out/${tag}_megakernel -f $P --no-flush > gpurun_out/${tag}_ncu_full.log 2>&1
  tail -3 gpurun_out/${tag}_ncu_full.log
  ;;
esac
