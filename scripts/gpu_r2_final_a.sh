# round-2 evidence, part A (1 GPU): bench line (S2 s = 1 headline with all side objects), GPU test log, smoke, reference
# arm, order sweep, shallow-water sizes, parity margins -- most important first
mkdir -p gpurun_out
export WXB_NO_REBUILD=1
nvidia-smi --query-gpu=name,memory.total --format=csv,noheader; nproc; free -g | head -2 | tail -1
( time timeout 1500 python bench.py ) > gpurun_out/r2_bench.json 2> gpurun_out/r2_bench.err; tail -4 gpurun_out/r2_bench.err
timeout 900 python -m pytest tests -m gpu -q -s 2>&1 | grep -E "error|passed|failed|drift|rhs error|perturbed" > gpurun_out/r2_gpu_tests.log; tail -2 gpurun_out/r2_gpu_tests.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r2_smoke.log 2>&1; tail -1 gpurun_out/r2_smoke.log
( time timeout 600 python bench.py --impl reference ) > gpurun_out/r2_bench_reference.json 2> gpurun_out/r2_bench_reference.err; tail -4 gpurun_out/r2_bench_reference.err
timeout 900 python scripts/sweep_orders.py > gpurun_out/r2_sweep_orders.jsonl 2> gpurun_out/r2_sweep_orders.err; tail -2 gpurun_out/r2_sweep_orders.err
timeout 600 python scripts/sw_bench.py galewsky 240,720,960 > gpurun_out/r2_sw_bench.jsonl 2>/dev/null
timeout 300 python scripts/sw_bench.py case6 960 >> gpurun_out/r2_sw_bench.jsonl 2>/dev/null
timeout 300 python scripts/gpu_errors.py > gpurun_out/r2_margins.log 2>&1
