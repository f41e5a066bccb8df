set -x
nvidia-smi -L
timeout 1500 python -m pytest tests/test_gpu_dp.py tests/test_gpu_invariance.py tests/test_gpu_module.py -q -m gpu 2>&1 | tail -8
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 30 --warmup 5 > gpurun_out/r2_bench_n2.json 2> gpurun_out/r2_bench_n2.err; tail -c 800 gpurun_out/r2_bench_n2.err
