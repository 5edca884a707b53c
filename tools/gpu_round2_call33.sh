set -x
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/r03k_pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r03k_pytest_gpu.log
tail -4 gpurun_out/r03k_pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r03k_smoke.log 2>&1; tail -1 gpurun_out/r03k_smoke.log
timeout 600 python bench.py > gpurun_out/r03k_bench_1gpu.json 2> gpurun_out/r03k_bench_1gpu.err; echo "bench rc=$?"
