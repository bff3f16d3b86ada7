# final evidence, part B (after part A's ncu captures): full GPU suite, the bench line with every side measurement and the
# CPU arm, the reference arm, launch list, parity margins, order sweep
set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q 2>&1 | tail -6 > gpurun_out/r1_gpu_tests.log; cat gpurun_out/r1_gpu_tests.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2 | tee gpurun_out/r1_smoke.log
timeout 900 python bench.py > gpurun_out/r1_bench.json 2> gpurun_out/r1_bench.err; echo rc=$?; tail -5 gpurun_out/r1_bench.err; cat gpurun_out/r1_bench.json
timeout 900 python bench.py --impl reference --steps 10 --warmup 2 > gpurun_out/r1_bench_reference.json 2>> gpurun_out/r1_bench.err; cat gpurun_out/r1_bench_reference.json
timeout 300 python scripts/gpu_errors.py > gpurun_out/r1_margins.log 2>&1; cat gpurun_out/r1_margins.log
ARGS="--steps 3 --warmup 3 --skip-check --skip-cpu --skip-sw --e2e-steps 3"
ncu --metrics gpu__time_duration.sum --clock-control none -c 160 --csv --log-file gpurun_out/r1_launches.csv python bench.py $ARGS > gpurun_out/ncu1.log 2>&1
timeout 600 python scripts/sweep_orders.py 2,3,4,5,6 > gpurun_out/r1_sweep_orders.jsonl 2>> gpurun_out/r1_bench.err; cat gpurun_out/r1_sweep_orders.jsonl
