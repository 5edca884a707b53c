set -x
mkdir -p gpurun_out
nvidia-smi -L
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 5 --warmup 3 > gpurun_out/r02k_bench_2gpu.json 2> gpurun_out/r02k_bench_2gpu.err; echo "bench rc=$?"
tail -3 gpurun_out/r02k_bench_2gpu.err
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 bench.py --impl reference --gpus 2 --steps 2 --warmup 1 > gpurun_out/r02k_bench_ref_2gpu.json 2> gpurun_out/r02k_bench_ref_2gpu.err; echo "ref rc=$?"
timeout 300 python tools/h2d_matrix.py --mb 512 > gpurun_out/r02k_h2d_matrix_2gpu.json 2>&1
cat gpurun_out/r02k_h2d_matrix_2gpu.json | head -40
# the multi-GPU call of the C ABI on two real devices
timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -q -k "multi or device_resident" > gpurun_out/r02k_pytest_multi.log 2>&1; tail -3 gpurun_out/r02k_pytest_multi.log
