set -x
mkdir -p gpurun_out
# 8 GPUs, config 2 only: share of a relaying rank's bytes that takes the relay (100 = all, -1 = balanced from the probe)
port=29540
for pct in ${PCTS:-100 -1 50}; do
port=$((port+1))
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port $port bench.py --gpus 8 --steps 5 --warmup 3 --no-config3 --c4-samples 0 --relay-percent $pct > gpurun_out/r02s_bench_8gpu_relay_pct$pct.json 2> gpurun_out/r02s_bench_8gpu_relay_pct$pct.err; echo "bench rc=$?"
done
