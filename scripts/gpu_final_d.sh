# final evidence, part D: the bench line with the final step kernels + one ncu capture of every step kernel
set -x
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_gpu_stepops.py tests/test_gpu_steps.py -m gpu -q 2>&1 | tail -3
timeout 900 python bench.py > gpurun_out/r1_bench.json 2> gpurun_out/r1_bench.err; echo rc=$?; tail -3 gpurun_out/r1_bench.err; tail -c 1500 gpurun_out/r1_bench.json
ARGS="--steps 3 --warmup 3 --skip-check --skip-cpu --skip-sw --e2e-steps 3"
WXB_STEPOPS_REPS=1 timeout 600 ncu --set full --clock-control none --import-source on -k regex:"expfilter|lincomb3|multidot|gs1sync|combine" -c 9 -f -o gpurun_out/r1_stepops python bench.py $ARGS > gpurun_out/ncu3.log 2>&1
tail -n 2 gpurun_out/ncu3.log | cut -c1-300
