TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
timeout 600 python -m pytest tests/test_gpu_parity.py tests/test_gpu_scale.py tests/test_gpu_round2.py -m gpu -q > gpurun_out/r02f_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r02f_pytest.log
timeout 300 python tools/dist_check.py > gpurun_out/r02_dist_check_n1.json 2> gpurun_out/r02f_dc1.err; echo "dc1 rc=$?"
timeout 300 $TR --master-port 29541 tools/dist_check.py > gpurun_out/r02_dist_check_n2.json 2> gpurun_out/r02f_dc2.err; echo "dc2 rc=$?"
cmp gpurun_out/r02_dist_check_n1.json gpurun_out/r02_dist_check_n2.json && echo "DIST CHECK: 1 GPU == 2 GPUs"
timeout 600 python tools/config_bench.py --configs 1,3,3r,4 > gpurun_out/r02f_configs_n1.jsonl 2> gpurun_out/r02f_cb1.err; echo "cb1 rc=$?"
timeout 600 $TR --master-port 29542 tools/config_bench.py --configs 1,3,3r,4 > gpurun_out/r02f_configs_n2.jsonl 2> gpurun_out/r02f_cb2.err; echo "cb2 rc=$?"
timeout 900 $TR --master-port 29543 bench.py --gpus 2 --steps 2 --warmup 2 --precision mixed > gpurun_out/r02f_bench_n2.json 2> gpurun_out/r02f_bench_n2.err; echo "bench rc=$?"
tail -4 gpurun_out/r02f_pytest.log; tail -3 gpurun_out/r02f_dc2.err; cat gpurun_out/r02f_configs_n1.jsonl gpurun_out/r02f_configs_n2.jsonl | cut -c1-500; tail -3 gpurun_out/r02f_cb2.err; cut -c1-300 gpurun_out/r02f_bench_n2.json; tail -3 gpurun_out/r02f_bench_n2.err
