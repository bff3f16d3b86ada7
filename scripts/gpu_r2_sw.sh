# shallow water: parity of the tile kernel, then both paths timed at scale (scripts/sw_bench.py)
WXB_NO_REBUILD=1 timeout 600 python -m pytest tests/test_gpu_parity.py tests/test_synthetic.py -m gpu -x -q -k "case6 or case5 or galewsky or sw or other_kernel" 2>&1 | tail -6
for p in tile split; do
echo "== WXB_SW_PATH=$p"
WXB_NO_REBUILD=1 WXB_SW_PATH=$p timeout 600 python scripts/sw_bench.py 2>&1 | tail -4 | cut -c1-400
done
