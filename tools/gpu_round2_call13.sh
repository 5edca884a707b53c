set -x
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -q -x > gpurun_out/r02m_pytest_parity_2gpu.log 2>&1; tail -3 gpurun_out/r02m_pytest_parity_2gpu.log
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 5 --warmup 3 --relay force:1 --no-config3 --c4-samples 0 > gpurun_out/r02m_bench_2gpu_relay_force1.json 2> gpurun_out/r02m_bench_2gpu_relay_force1.err; echo "bench rc=$?"
tail -3 gpurun_out/r02m_bench_2gpu_relay_force1.err
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29513 bench.py --gpus 2 --steps 5 --warmup 3 --no-config3 --c4-samples 0 > gpurun_out/r02m_bench_2gpu_auto.json 2> gpurun_out/r02m_bench_2gpu_auto.err; echo "bench rc=$?"
