#!/bin/bash
# 8-GPU record of round 2 (run on the GPU box): selections vs the committed 1-GPU record, configs 1/3/3r/4 with stage
# breakdowns, the strong-scaling bench, config 5.
N=${1:-8}
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
timeout 400 $TR --master-port 29551 tools/dist_check.py 2> gpurun_out/r02h_dc.err | grep -v '^NCCL' > gpurun_out/r02_dist_check_n$N.json; echo "dc rc=$?"
cmp gpurun_out/r02_dist_check_n$N.json profiles/r02_dist_check_n1.json && echo "DIST CHECK: $N GPUs == 1 GPU"
timeout 600 $TR --master-port 29552 tools/config_bench.py --configs 1,3,3r,4 2> gpurun_out/r02h_cb.err | grep '^{' > gpurun_out/r02_configs_n$N.jsonl; echo "cb rc=$?"
timeout 900 $TR --master-port 29553 bench.py --gpus $N --steps 3 --warmup 3 2> gpurun_out/r02h_bench.err | grep '^{' > gpurun_out/r02_bench_n${N}_mixed.json; echo "bench rc=$?"
timeout 600 $TR --master-port 29554 tools/config5.py --pool 2048 --rounds 3 --iters 60 --out gpurun_out/r02_config5_n$N.jsonl > gpurun_out/r02h_c5.log 2>&1; echo "c5 rc=$?"
cat gpurun_out/r02_configs_n$N.jsonl | cut -c1-420; cut -c1-300 gpurun_out/r02_bench_n${N}_mixed.json; tail -4 gpurun_out/r02h_c5.log | cut -c1-300; tail -3 gpurun_out/r02h_dc.err gpurun_out/r02h_cb.err gpurun_out/r02h_bench.err | cut -c1-300
