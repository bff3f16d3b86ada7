# round 2 baseline: the round-1 kernels at SURVEY 8(d) S2 s = 1 (24x96x96 elements/panel, 85 M nodes) and s = 1/2
nvidia-smi --query-gpu=name,memory.total --format=csv,noheader
free -g | head -2
nproc
for cfg in "12 48" "24 96"; do
set -- $cfg
echo "== nz $1 nh $2"
/usr/bin/time -v timeout 900 python bench.py --nz $1 --nh $2 --steps 20 --warmup 5 --skip-cpu --skip-sw --skip-stepops --e2e-steps 3 > gpurun_out/r2_base_$1_$2.json 2> gpurun_out/r2_base_$1_$2.err
grep -E "Maximum resident|Elapsed" gpurun_out/r2_base_$1_$2.err
tail -1 gpurun_out/r2_base_$1_$2.json | python -c "
import json,sys
d=json.loads(sys.stdin.read())
print('value %.3e  ms/step %.3f  e2e %.3e setup %.1f' % (d['value'], d['ms_per_step'], d['e2e']['value'], d['setup_s']))
for k in ('roofline','roofline_jvp'):
    r=d[k]; print(k, 'frac %.3f' % r['frac'], r['launch_ms'])
print(d['parity'], d['clocks'])
"
done
