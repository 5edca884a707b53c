set -x
mkdir -p gpurun_out
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29581 bench.py --gpus 8 > gpurun_out/r03c_bench_8gpu_default.json 2> gpurun_out/r03c_bench_8gpu_default.err; echo "bench rc=$?"
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29582 bench.py --gpus 4 > gpurun_out/r03c_bench_4gpu_default.json 2> gpurun_out/r03c_bench_4gpu_default.err; echo "bench rc=$?"
