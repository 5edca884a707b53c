set -x
mkdir -p gpurun_out
( time timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29571 bench.py --gpus 2 > gpurun_out/r02y_bench_2gpu.json 2> gpurun_out/r02y_bench_2gpu.err ) 2> gpurun_out/r02y_bench_2gpu_time.txt; tail -3 gpurun_out/r02y_bench_2gpu_time.txt
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29572 bench.py --gpus 2 --impl reference --steps 2 --warmup 1 > gpurun_out/r02y_bench_2gpu_ref.json 2> gpurun_out/r02y_bench_2gpu_ref.err; tail -c 600 gpurun_out/r02y_bench_2gpu_ref.json
