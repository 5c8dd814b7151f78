set -x
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1"
timeout 400 python -m pytest tests/test_multigpu.py -q -m gpu -x -s 2>&1 | grep -v "^W1\|OMP_NUM" | tail -15
timeout 240 $TR --master-port 29541 bench.py --gpus 8 --steps 200 --warmup 20 2>/dev/null | tail -1 > gpurun_out/bench_ours_n8.json; cut -c1-330 gpurun_out/bench_ours_n8.json
B200DDP_ARENA=symm_mem timeout 240 $TR --master-port 29542 benchmarks/allreduce_sweep.py --graph --max-mb 256 --out gpurun_out/allreduce_sweep_n8_graph_nvls.json 2>&1 | grep "^{" | tail -30 | cut -c1-260
timeout 240 $TR --master-port 29543 benchmarks/bench_models.py --model gpt2-medium --impl ours --steps 8 --warmup 3 2>&1 | grep "^{" | tail -1 | tee gpurun_out/gpt2_medium_ours_n8.json
timeout 240 $TR --master-port 29544 benchmarks/bench_models.py --model gpt2-medium --impl stock --steps 8 --warmup 3 2>&1 | grep "^{" | tail -1 | tee gpurun_out/gpt2_medium_stock_n8.json
timeout 240 $TR --master-port 29545 benchmarks/bench_models.py --model resnet50 --impl ours --steps 20 --warmup 5 2>&1 | grep "^{" | tail -1 | tee gpurun_out/resnet50_ours_n8.json
timeout 240 $TR --master-port 29546 benchmarks/bench_models.py --model resnet50 --impl stock --steps 20 --warmup 5 2>&1 | grep "^{" | tail -1 | tee gpurun_out/resnet50_stock_n8.json
