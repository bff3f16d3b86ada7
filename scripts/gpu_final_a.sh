# final evidence, part A: GPU suite (both Euler kernel variants), launch list, ncu full capture of the Euler kernels at
# the bench size and of the per-step kernels
set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q 2>&1 | tail -6 > gpurun_out/r1_gpu_tests.log; cat gpurun_out/r1_gpu_tests.log
WXB_KERNEL_OPTS=4 timeout 600 python -m pytest tests/test_gpu_parity.py tests/test_gpu_drift.py -m gpu -q 2>&1 | tail -3 > gpurun_out/r1_gpu_tests_staged.log; cat gpurun_out/r1_gpu_tests_staged.log
timeout 300 python scripts/gpu_errors.py > gpurun_out/r1_margins.log 2>&1; cat gpurun_out/r1_margins.log
ARGS="--steps 3 --warmup 3 --skip-check --skip-cpu --skip-sw --e2e-steps 3"
timeout 600 python bench.py $ARGS > gpurun_out/plain.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 120 --csv --log-file gpurun_out/r1_launches.csv python bench.py $ARGS > gpurun_out/ncu1.log 2>&1
ARGS="$ARGS --skip-stepops"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:euler_ -s 12 -c 4 -f -o gpurun_out/r1_final python bench.py $ARGS > gpurun_out/ncu2.log 2>&1
ARGS="--steps 3 --warmup 3 --skip-check --skip-cpu --skip-sw --e2e-steps 3"
WXB_STEPOPS_REPS=1 timeout 900 ncu --set full --clock-control none --import-source on -k regex:"expfilter|lincomb3|multidot|gs1sync|combine" -c 8 -f -o gpurun_out/r1_stepops python bench.py $ARGS > gpurun_out/ncu3.log 2>&1
tail -n 3 gpurun_out/ncu1.log gpurun_out/ncu2.log gpurun_out/ncu3.log
