# both Euler paths at the headline size (S2 s = 1) and at s = 1/2, current kernels
for path in sweep split; do
for cfg in "24 96" "12 48"; do
set -- $cfg
echo "== path $path nz $1 nh $2"
WXB_NO_REBUILD=1 WXB_EULER_PATH=$path timeout 900 python bench.py --nz $1 --nh $2 --steps 20 --warmup 5 --repeats 3 --skip-sweep --skip-cpu --skip-sw --skip-stepops --e2e-steps 3 2> gpurun_out/r2p_${path}_$1.err | tail -1 > gpurun_out/r2p_${path}_$1.json
tail -2 gpurun_out/r2p_${path}_$1.err
python - <<PY
import json
d=json.load(open("gpurun_out/r2p_${path}_$1.json"))
print("value %.3e ms/step %.3f e2e %.3e setup %.1f" % (d["value"], d["ms_per_step"], d["e2e"]["value"], d["setup_s"]))
for k in ("roofline","roofline_jvp"):
    r=d[k]; print(k, "burst %.3f (%.3f ms) sustained %.3f" % (r["frac"], r["launch_ms"]["evaluation"], r["sustained"]["frac"]))
print(d["parity"]["ok"], d["parity"]["rhs_rel_err"], d["clocks"], d["sustained"]["clocks"])
PY
done
done
