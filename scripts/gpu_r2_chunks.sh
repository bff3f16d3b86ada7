# end-to-end step (host arrays) against the number of level chunks of the host pipeline, S2 s = 1, 1 GPU
export WXB_NO_REBUILD=1
ARGS="--steps 3 --warmup 3 --repeats 1 --sustain-s 0 --skip-check --skip-sweep --skip-strong --skip-cpu --skip-sw --skip-stepops --e2e-steps 6"
for c in 8 12 24; do
  WXB_HOST_CHUNKS=$c timeout 600 python bench.py $ARGS 2>/dev/null | tail -1 | python -c "
import json,sys
d=json.loads(sys.stdin.read()); e=d['e2e']
print('chunks $c: e2e %.1f ms/step, %.3e, calls %s' % (e['ms_per_step'], e['value'], {k: round(v,1) for k,v in e['host_call_ms'].items() if k!='what'}))"
done
