set -x
timeout 900 python -m pytest tests/test_gpu_module.py tests/test_gpu_dp.py -q -m gpu -x 2>&1 | tail -4
python profiles/experiments/small_batch_probe.py 2>&1 | tail -1
python profiles/train_quality.py 18 > gpurun_out/r2_train_quality.json 2> gpurun_out/r2_train_quality.err; cat gpurun_out/r2_train_quality.err | tail -20
