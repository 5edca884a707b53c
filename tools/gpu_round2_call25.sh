set -x
mkdir -p gpurun_out
python tools/cubin_key.py > gpurun_out/r03d_cubin_key.txt
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_zz_no_alloc.py -m gpu -x -q > gpurun_out/r03d_pytest_parity.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r03d_pytest_parity.log
tail -4 gpurun_out/r03d_pytest_parity.log
timeout 600 python bench.py --no-cpu-baseline --c4-samples 0 > gpurun_out/r03d_bench_1gpu.json 2> gpurun_out/r03d_bench_1gpu.err; echo "bench rc=$?"
timeout 600 python bench.py --no-cpu-baseline --c4-samples 0 > gpurun_out/r03d_bench_1gpu_b.json 2> gpurun_out/r03d_bench_1gpu_b.err; echo "bench rc=$?"
