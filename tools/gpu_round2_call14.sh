set -x
mkdir -p gpurun_out
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29521 bench.py --gpus 8 --steps 5 --warmup 3 --no-config3 --c4-samples 0 > gpurun_out/r02n_bench_8gpu_relay_auto.json 2> gpurun_out/r02n_bench_8gpu_relay_auto.err; echo "bench rc=$?"
tail -3 gpurun_out/r02n_bench_8gpu_relay_auto.err
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29523 bench.py --gpus 8 --steps 5 --warmup 3 --no-config3 --c4-samples 0 --relay off > gpurun_out/r02n_bench_8gpu_relay_off.json 2> gpurun_out/r02n_bench_8gpu_relay_off.err; echo "bench rc=$?"
