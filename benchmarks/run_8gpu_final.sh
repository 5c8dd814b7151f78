#!/bin/bash
# Final 8-GPU pass of a round (run through `gpurun --gpus 8`): headline at 8 and 4 ranks with the driver's flags,
# the two extra workloads at 8 ranks, the multi-GPU end-to-end tests and the per-kernel DDP timeline.
set -x
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
timeout 300 python bench.py --gpus 8 --steps 20 --warmup 5 > gpurun_out/f8_ours_n8.json 2> gpurun_out/f8_ours_n8.err
timeout 300 python bench.py --gpus 8 --steps 200 --warmup 20 > gpurun_out/f8_ours_n8_long.json 2>/dev/null
timeout 300 python bench.py --gpus 4 --steps 20 --warmup 5 > gpurun_out/f8_ours_n4.json 2>/dev/null
timeout 300 python bench.py --gpus 8 --model gpt2-medium --steps 10 --warmup 3 > gpurun_out/f8_gpt2_ours_n8.json 2> gpurun_out/f8_gpt2_ours_n8.err
timeout 300 python bench.py --gpus 8 --model resnet50 --steps 20 --warmup 5 > gpurun_out/f8_resnet50_ours_n8.json 2> gpurun_out/f8_resnet50_ours_n8.err
(timeout 300 python -m pytest tests/test_multigpu.py -q -m gpu -s -k "collectives or end_to_end or hooks" 2>&1 \
   | grep -v "^W1\|OMP_NUM\|^\*\*\*" | tail -8) > gpurun_out/f8_pytest.log
B200DDP_SEQ_OUT=gpurun_out/f8_seq_n8.txt timeout 200 $TR --nproc-per-node 8 --master-port 29545 benchmarks/timeline_ddp.py \
   > gpurun_out/f8_timeline_n8.txt 2>&1
tail -4 gpurun_out/f8_pytest.log
for f in f8_ours_n8 f8_ours_n8_long f8_ours_n4 f8_gpt2_ours_n8 f8_resnet50_ours_n8; do python - "$f" <<'PY'
import json, sys
f = sys.argv[1]
try:
    d = json.load(open(f"gpurun_out/{f}.json"))
    b = d.get("step_breakdown") or {}
    print(f, d.get("value"), d.get("ms_per_step"), (d.get("e2e") or {}).get("ms_per_step"),
          {k: round(v, 1) for k, v in b.items() if isinstance(v, float)}, d.get("unavailable"))
except Exception as e:
    print(f, "unreadable", e)
PY
done
