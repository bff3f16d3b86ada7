# round-1 evidence run: GPU tests (both kernel variants), bench, ncu launch list + full capture of both kernels
set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q 2>&1 | tail -15 > gpurun_out/r1_gpu_tests.log; cat gpurun_out/r1_gpu_tests.log
WXB_KERNEL_OPTS=4 timeout 900 python -m pytest tests -m gpu -q 2>&1 | tail -5 > gpurun_out/r1_gpu_tests_staged.log; cat gpurun_out/r1_gpu_tests_staged.log
timeout 900 python bench.py > gpurun_out/r1_bench.json 2> gpurun_out/r1_bench.err; echo rc=$?; tail -5 gpurun_out/r1_bench.err; cat gpurun_out/r1_bench.json
timeout 900 python bench.py --impl reference --steps 10 --warmup 2 > gpurun_out/r1_bench_reference.json 2>> gpurun_out/r1_bench.err; cat gpurun_out/r1_bench_reference.json
ARGS="--steps 3 --warmup 3 --nh 24 --skip-check --skip-cpu --e2e-steps 3"
timeout 600 python bench.py $ARGS > gpurun_out/plain.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 60 --csv --log-file gpurun_out/r1_launches.csv python bench.py $ARGS > gpurun_out/ncu1.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:euler_ -s 8 -c 4 -f -o gpurun_out/r1_final python bench.py $ARGS > gpurun_out/ncu2.log 2>&1
tail -n 3 gpurun_out/ncu1.log gpurun_out/ncu2.log
