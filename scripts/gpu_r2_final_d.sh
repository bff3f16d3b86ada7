# round-2 evidence, part D (1 GPU): the size rule of the default Euler path (small handles -> element-parallel pair):
# GPU tests, smoke, the order sweep, and the bench's size sweep only
mkdir -p gpurun_out
export WXB_NO_REBUILD=1
timeout 900 python -m pytest tests -m gpu -q -s 2>&1 | grep -E "error|passed|failed|drift|rhs error|perturbed" > gpurun_out/r2_gpu_tests.log; tail -2 gpurun_out/r2_gpu_tests.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r2_smoke.log 2>&1; tail -1 gpurun_out/r2_smoke.log
timeout 900 python scripts/sweep_orders.py > gpurun_out/r2_sweep_orders.jsonl 2> gpurun_out/r2_sweep_orders.err; tail -2 gpurun_out/r2_sweep_orders.err
python - <<'PY'
import json
for l in open("gpurun_out/r2_sweep_orders.jsonl"):
    d = json.loads(l); print(d["n"], d["shape"], d["kernel_path"], round(d["rhs_ms"], 3), round(d["rhs_frac_hbm"], 3), round(d["jvp_ms"], 3), round(d["jvp_frac_hbm"], 3))
PY
( time timeout 900 python bench.py --skip-strong ) > gpurun_out/r2_bench.json 2> gpurun_out/r2_bench.err; tail -4 gpurun_out/r2_bench.err
