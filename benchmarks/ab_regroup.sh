#!/usr/bin/env bash
# regroup kernel (variant 1) sweep: refill threshold x triangle-batch threshold x resident CTAs/SM, against the megakernel (variant 0)
mkdir -p gpurun_out
ms() { python -c "import json,sys; d=json.loads(sys.stdin.read()); print(round(d['ms_per_step'],2))"; }
echo "parity: $(timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k 'regroup' 2>&1 | tail -1)" | tee -a gpurun_out/ab_regroup.txt
echo "variant0 ms=$(timeout 300 python bench.py --steps 8 --warmup 3 --no-cpu-baseline --no-e2e 2>/dev/null | ms)" | tee -a gpurun_out/ab_regroup.txt
for mb in ${MBS:-3 4}; do for r in ${REFILLS:-8 16}; do for lb in ${BATCHES:-1 6 12}; do
  echo "variant1 minblocks=$mb refill=$r leaf_batch=$lb ms=$(B200RT_RG_MINBLOCKS=$mb B200RT_REFILL=$r B200RT_LEAF_BATCH=$lb timeout 300 python bench.py --variant 1 --steps 8 --warmup 3 --no-cpu-baseline --no-e2e 2>/dev/null | ms)" | tee -a gpurun_out/ab_regroup.txt
done; done; done
