set -x
mkdir -p gpurun_out
# 8 GPUs, the default line (relay share tuned over two untimed steps), then config 2 alone twice for the spread of the tuned share
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29551 bench.py --gpus 8 > gpurun_out/r02t_bench_8gpu_default.json 2> gpurun_out/r02t_bench_8gpu_default.err; echo "bench rc=$?"
for k in 1 2; do
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 2956$k bench.py --gpus 8 --steps 5 --warmup 3 --no-config3 --c4-samples 0 > gpurun_out/r02t_bench_8gpu_tuned_run$k.json 2> gpurun_out/r02t_bench_8gpu_tuned_run$k.err; echo "bench rc=$?"
done
