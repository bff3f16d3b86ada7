# end-to-end (host arrays) step time against the number of level chunks of the host pipeline
for c in ${@:-1 3 6 12}; do
echo "== WXB_HOST_CHUNKS=$c"
WXB_HOST_CHUNKS=$c timeout 600 python bench.py --steps 10 --warmup 3 --skip-cpu --skip-sw --skip-stepops --e2e-steps 8 2>&1 | tail -1 | python -c "
import json,sys
d=json.loads(sys.stdin.read())
e=d['e2e']; print('e2e %.3e  %.2f ms/step  h2d %.0f MB d2h %.0f MB | device %.3e' % (e['value'], e['ms_per_step'], e['h2d_bytes_per_step']/1e6, e['d2h_bytes_per_step']/1e6, d['value']))
"
done
