# final evidence, 8 ranks: the weak-scaling bench line (decomposition invariance at 8 ranks: profiles/r01_multi_check_n8.log)
mkdir -p gpurun_out
timeout 100 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus 8 --steps 20 --warmup 5 --e2e-steps 5 2> gpurun_out/r1_bench_8.err | tail -1 > gpurun_out/r1_bench_8.json
tail -2 gpurun_out/r1_bench_8.err; cut -c1-400 gpurun_out/r1_bench_8.json
