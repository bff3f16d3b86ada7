# final evidence, part E: step-kernel tests with the reduction kernels, ncu full capture of the final Euler kernels
set -x
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_gpu_stepops.py tests/test_gpu_steps.py -m gpu -q -s 2>&1 | grep -E "drift|passed|failed|Error" | tee gpurun_out/r1_gpu_steps.log
ARGS="--steps 3 --warmup 3 --skip-check --skip-cpu --skip-sw --skip-stepops --e2e-steps 3"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:euler_ -s 12 -c 4 -f -o gpurun_out/r1_final3 python bench.py $ARGS > gpurun_out/ncu2.log 2>&1
tail -n 2 gpurun_out/ncu2.log | cut -c1-200
