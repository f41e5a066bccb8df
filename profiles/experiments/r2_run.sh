set -x
python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-profile --no-sub > gpurun_out/plain.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -s 40 -c 200 --csv --log-file gpurun_out/r2_launches_infer.csv python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-profile --no-sub > gpurun_out/ncu1.log 2>&1
python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-profile --no-sub > gpurun_out/plain.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:k_tc_fwd -s 4 -c 1 -o gpurun_out/r2_tc_fwd_final -f python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-profile --no-sub > gpurun_out/ncu2.log 2>&1
python bench.py --mode train --steps 3 --warmup 3 --no-cpu-baseline --no-profile > gpurun_out/plain.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:k_tc_bwd -s 4 -c 1 -o gpurun_out/r2_tc_bwd -f python bench.py --mode train --steps 3 --warmup 3 --no-cpu-baseline --no-profile > gpurun_out/ncu3.log 2>&1
python bench.py --mode train --steps 3 --warmup 3 --no-cpu-baseline --no-profile > gpurun_out/plain.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -s 60 -c 200 --csv --log-file gpurun_out/r2_launches_train.csv python bench.py --mode train --steps 3 --warmup 3 --no-cpu-baseline --no-profile > gpurun_out/ncu4.log 2>&1
tail -2 gpurun_out/ncu2.log | cut -c1-150; tail -2 gpurun_out/ncu3.log | cut -c1-150
