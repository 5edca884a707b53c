set -x
mkdir -p gpurun_out
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29597 bench.py --gpus 8 --steps 5 --warmup 3 --no-config3 --c4-samples 0 > gpurun_out/r03j_bench_8gpu_check.json 2> gpurun_out/r03j_bench_8gpu_check.err; echo "bench rc=$?"
tail -3 gpurun_out/r03j_bench_8gpu_check.err
