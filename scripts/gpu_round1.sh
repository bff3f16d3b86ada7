# round-1 evidence run: GPU tests (both Euler kernel variants), bench, reference arm, shallow-water sweep, order sweep (quick),
# ncu launch list + full captures (Euler at the bench size, shallow water)
set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q 2>&1 | tail -15 > gpurun_out/r1_gpu_tests.log; cat gpurun_out/r1_gpu_tests.log
WXB_KERNEL_OPTS=4 timeout 900 python -m pytest tests -m gpu -q 2>&1 | tail -5 > gpurun_out/r1_gpu_tests_staged.log; cat gpurun_out/r1_gpu_tests_staged.log
timeout 900 python bench.py > gpurun_out/r1_bench.json 2> gpurun_out/r1_bench.err; echo rc=$?; tail -5 gpurun_out/r1_bench.err; cat gpurun_out/r1_bench.json
timeout 900 python bench.py --impl reference --steps 10 --warmup 2 > gpurun_out/r1_bench_reference.json 2>> gpurun_out/r1_bench.err; cat gpurun_out/r1_bench_reference.json
timeout 600 python scripts/sw_bench.py galewsky 240,960 > gpurun_out/r1_sw_bench.jsonl 2>> gpurun_out/r1_bench.err; cat gpurun_out/r1_sw_bench.jsonl
timeout 600 python scripts/sw_bench.py case6 960 >> gpurun_out/r1_sw_bench.jsonl 2>> gpurun_out/r1_bench.err; tail -1 gpurun_out/r1_sw_bench.jsonl
timeout 900 python scripts/sweep_orders.py 2,3,4,5,6 > gpurun_out/r1_sweep_orders.jsonl 2>> gpurun_out/r1_bench.err; cat gpurun_out/r1_sweep_orders.jsonl
python scripts/gpu_errors.py > gpurun_out/r1_margins.log 2>&1; cat gpurun_out/r1_margins.log
ARGS="--steps 3 --warmup 3 --skip-check --skip-cpu --e2e-steps 3"
timeout 600 python bench.py $ARGS > gpurun_out/plain.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 80 --csv --log-file gpurun_out/r1_launches.csv python bench.py $ARGS > gpurun_out/ncu1.log 2>&1
ARGS="$ARGS --skip-sw"
timeout 1500 ncu --set full --clock-control none --import-source on -k regex:euler_ -s 12 -c 4 -f -o gpurun_out/r1_final python bench.py $ARGS > gpurun_out/ncu2.log 2>&1
timeout 300 python scripts/sw_bench.py galewsky 480 > gpurun_out/sw_plain.log 2>&1 && \
timeout 900 ncu --set full --clock-control none --import-source on -k regex:sw_ -s 6 -c 4 -f -o gpurun_out/r1_final_sw python scripts/sw_bench.py galewsky 480 > gpurun_out/ncu3.log 2>&1
tail -n 3 gpurun_out/ncu1.log gpurun_out/ncu2.log gpurun_out/ncu3.log
