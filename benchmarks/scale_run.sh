#!/usr/bin/env bash
# scale_run.sh <tag> <N>...: bench.py under torchrun at each N (one node), JSON lines to gpurun_out/<tag>_bench_<N>gpu.json
tag=$1; shift
mkdir -p gpurun_out
for n in "$@"; do
  if [ "$n" = 1 ]; then python bench.py --gpus 1 --no-cpu-baseline --no-configs > gpurun_out/${tag}_bench_1gpu.json 2> gpurun_out/${tag}_bench_1gpu.err
  else python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29500 + n)) bench.py --gpus $n > gpurun_out/${tag}_bench_${n}gpu.json 2> gpurun_out/${tag}_bench_${n}gpu.err; fi
  python - <<PY
import json
try:
    d = json.loads(open("gpurun_out/${tag}_bench_${n}gpu.json").read())
    print($n, "value", round(d["value"]), "ms", round(d["ms_per_step"], 2), "e2e", round(d["e2e"]["value"]), "strong ms", round(d["strong"]["ms"], 1), "reduce", round(d["strong"]["reduce_ms"], 3), "frac", d["roofline"]["frac"])
except Exception as e:
    print($n, "failed", e); print(open("gpurun_out/${tag}_bench_${n}gpu.err").read()[-1500:])
PY
done
