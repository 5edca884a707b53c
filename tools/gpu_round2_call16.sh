set -x
mkdir -p gpurun_out
for pct in -1 35 50; do
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 2953$((pct+2)) bench.py --gpus 8 --steps 5 --warmup 3 --no-config3 --c4-samples 0 --relay-percent $pct > gpurun_out/r02q_bench_8gpu_relay_pct$pct.json 2> gpurun_out/r02q_bench_8gpu_relay_pct$pct.err; echo "bench rc=$?"
done
