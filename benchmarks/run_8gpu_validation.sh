#!/bin/bash
# 8-GPU validation pass (run through `gpurun --gpus 8`): multi-GPU tests on 8 ranks, the headline bench at 8 / 4
# ranks for both arms, the all-reduce sweep against NCCL at 8 and 4 ranks.  Everything lands in gpurun_out/.
set -x
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
(timeout 420 python -m pytest tests/test_multigpu.py -q -m gpu -s -k "collectives or end_to_end or nvls or dead_rank" 2>&1 \
   | grep -v "^W1\|OMP_NUM\|^\*\*\*" | tail -25) > gpurun_out/v8_pytest.log
timeout 300 python bench.py --impl reference --gpus 8 --steps 20 --warmup 5 > gpurun_out/v8_ref_n8.json 2> gpurun_out/v8_ref_n8.err
timeout 300 python bench.py --gpus 8 --steps 20 --warmup 5 > gpurun_out/v8_ours_n8.json 2> gpurun_out/v8_ours_n8.err
timeout 300 python bench.py --gpus 8 --steps 200 --warmup 20 > gpurun_out/v8_ours_n8_long.json 2>/dev/null
timeout 300 python bench.py --gpus 4 --steps 20 --warmup 5 > gpurun_out/v8_ours_n4.json 2>/dev/null
timeout 300 python bench.py --impl reference --gpus 4 --steps 20 --warmup 5 > gpurun_out/v8_ref_n4.json 2>/dev/null
timeout 400 $TR --nproc-per-node 8 --master-port 29542 benchmarks/allreduce_sweep.py --graph --max-mb 1024 \
   --out gpurun_out/allreduce_sweep_n8.json > gpurun_out/v8_sweep8.log 2>&1
timeout 300 $TR --nproc-per-node 4 --master-port 29543 benchmarks/allreduce_sweep.py --graph --max-mb 1024 \
   --out gpurun_out/allreduce_sweep_n4.json > gpurun_out/v8_sweep4.log 2>&1
B200DDP_NVLS_BLOCKS=48 timeout 200 $TR --nproc-per-node 8 --master-port 29544 benchmarks/allreduce_sweep.py --graph --max-mb 256 \
   --sizes buckets --no-nccl --out gpurun_out/allreduce_sweep_n8_nvls48.json > gpurun_out/v8_sweep8_nvls48.log 2>&1
B200DDP_SEQ_OUT=gpurun_out/v8_seq_n8.txt timeout 200 $TR --nproc-per-node 8 --master-port 29545 benchmarks/timeline_ddp.py \
   > gpurun_out/v8_timeline_n8.txt 2>&1
tail -6 gpurun_out/v8_pytest.log
for f in v8_ref_n8 v8_ours_n8 v8_ours_n8_long v8_ours_n4 v8_ref_n4; do python - "$f" <<'PY'
import json, sys
f = sys.argv[1]
try:
    d = json.load(open(f"gpurun_out/{f}.json"))
    print(f, d.get("value"), d.get("ms_per_step"), (d.get("e2e") or {}).get("value"), d.get("allreduce"), d.get("step_breakdown"), d.get("unavailable"))
except Exception as e:
    print(f, "unreadable", e)
PY
done
grep "^{" gpurun_out/v8_sweep8.log | cut -c1-330 | tail -32
