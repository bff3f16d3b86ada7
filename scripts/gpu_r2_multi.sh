# multi-GPU check + bench: bash scripts/gpu_r2_multi.sh N [bench args]
N=${1:-2}; shift
mkdir -p gpurun_out
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 tests/multi_rank_check.py 2>&1 | grep -v Warning | tail -14 | tee gpurun_out/r2_multi_check_$N.log
timeout 1500 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus $N --steps 20 --warmup 5 "$@" 2> gpurun_out/r2_bench_$N.err | tail -1 > gpurun_out/r2_bench_$N.json
tail -5 gpurun_out/r2_bench_$N.err
python - <<PY
import json
d=json.load(open("gpurun_out/r2_bench_$N.json"))
print("N=%d value %.3e ms/step %.3f e2e %.3e (%.1f ms)" % (d["n_gpus"], d["value"], d["ms_per_step"], d["e2e"]["value"], d["e2e"]["ms_per_step"]))
for k in ("roofline","roofline_jvp"): print(k, "%.3f" % d[k]["frac"], d[k]["launch_ms"], d[k].get("sustained"))
print(d["parity"]["ok"], d["parity"]["rhs_rel_err"], d["clocks"], d["config"]["workload"])
print("strong", d.get("strong"))
PY
