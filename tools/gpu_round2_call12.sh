set -x
mkdir -p gpurun_out
nvidia-smi -L | head -8
bash tools/host_topology.sh > gpurun_out/r02l_host_topology_8gpu.txt 2>&1
timeout 300 python tools/h2d_matrix.py --mb 1024 --reps 6 > gpurun_out/r02l_h2d_matrix_8gpu.json 2> gpurun_out/r02l_h2d_matrix_8gpu.err
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29521 bench.py --gpus 8 --steps 5 --warmup 3 > gpurun_out/r02l_bench_8gpu.json 2> gpurun_out/r02l_bench_8gpu.err; echo "bench rc=$?"
tail -3 gpurun_out/r02l_bench_8gpu.err
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29522 bench.py --gpus 4 --steps 5 --warmup 3 > gpurun_out/r02l_bench_4gpu.json 2> gpurun_out/r02l_bench_4gpu.err; echo "bench rc=$?"
