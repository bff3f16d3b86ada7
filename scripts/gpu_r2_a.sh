# round 2, first look at the column-sweep kernel: parity (both Euler paths), then both paths timed at s = 1/2 and s = 1
nvidia-smi --query-gpu=name,memory.total --format=csv,noheader; free -g | head -2 | tail -1; nproc
timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -q -x 2>&1 | tail -15
for path in sweep split; do
for cfg in "12 48" "24 96"; do
set -- $cfg
echo "== path $path nz $1 nh $2"
WXB_EULER_PATH=$path timeout 900 python bench.py --nz $1 --nh $2 --steps 20 --warmup 5 --skip-cpu --skip-sw --skip-stepops --e2e-steps 3 > gpurun_out/r2a_${path}_$1_$2.json 2> gpurun_out/r2a_${path}_$1_$2.err
tail -3 gpurun_out/r2a_${path}_$1_$2.err
tail -1 gpurun_out/r2a_${path}_$1_$2.json | python -c "
import json,sys
d=json.loads(sys.stdin.read())
print('value %.3e  ms/step %.3f  e2e %.3e setup %.1f' % (d['value'], d['ms_per_step'], d['e2e']['value'], d['setup_s']))
for k in ('roofline','roofline_jvp'):
    r=d[k]; print(k, 'frac %.3f' % r['frac'], r['launch_ms'])
print(d['parity'], d['clocks'])
"
done
done
