set -x
mkdir -p gpurun_out
python tools/cubin_key.py > gpurun_out/r02h_cubin_key.txt
timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -x -q > gpurun_out/r02h_pytest_parity.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r02h_pytest_parity.log
tail -3 gpurun_out/r02h_pytest_parity.log
timeout 600 python bench.py > gpurun_out/r02h_bench_1gpu.json 2> gpurun_out/r02h_bench_1gpu.err; echo "bench rc=$?"
tail -3 gpurun_out/r02h_bench_1gpu.err
