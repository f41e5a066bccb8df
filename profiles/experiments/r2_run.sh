set -x
B="python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-profile --no-sub"
$B > gpurun_out/plain.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -s 30 -c 120 --csv --log-file gpurun_out/r2b_launches_infer.csv $B > gpurun_out/ncu1.log 2>&1
$B > gpurun_out/plain.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:k_tc_fwd -s 4 -c 1 -o gpurun_out/r2b_tc_fwd_coo -f $B > gpurun_out/ncu2.log 2>&1
T="python bench.py --mode train --steps 3 --warmup 3 --no-cpu-baseline --no-profile"
$T > gpurun_out/plain.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -s 60 -c 200 --csv --log-file gpurun_out/r2b_launches_train.csv $T > gpurun_out/ncu4.log 2>&1
PYTHONPATH=. python profiles/experiments/ensemble_rate.py 4096 > gpurun_out/plain.log 2>&1 && \
PYTHONPATH=. ncu --set full --clock-control none --import-source on -k regex:k_tc_fwd_ens -s 12 -c 1 -o gpurun_out/r2b_tc_fwd_ens -f python profiles/experiments/ensemble_rate.py 4096 > gpurun_out/ncu5.log 2>&1
tail -2 gpurun_out/ncu2.log | cut -c1-150; tail -2 gpurun_out/ncu5.log | cut -c1-150
