# final evidence, part C: N ranks -- decomposition invariance (RHS/JVP bit-identical to one rank, integrators with the
# Krylov vectors split over the ranks) and the weak-scaling bench line: bash scripts/gpu_final_c.sh N
N=${1:-2}
mkdir -p gpurun_out
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 tests/multi_rank_check.py 2>&1 | grep -v Warning | tail -14 | tee gpurun_out/r1_multi_check_$N.log
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus $N --steps 20 --warmup 5 2> gpurun_out/r1_bench_$N.err | tail -1 > gpurun_out/r1_bench_$N.json
tail -3 gpurun_out/r1_bench_$N.err
python - <<PY
import json
d=json.load(open("gpurun_out/r1_bench_$N.json"))
print("N=%d value %.3e ms/step %.3f e2e %.3e" % (d["n_gpus"], d["value"], d["ms_per_step"], d["e2e"]["value"]))
for k in ("roofline","roofline_jvp"): print(k, "%.3f" % d[k]["frac"], d[k]["launch_ms"])
print(d["parity"]["ok"], d["parity"]["rhs_rel_err"], d["clocks"], d["config"]["workload"])
PY
