set -x
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q --durations=6 > gpurun_out/r03f_pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r03f_pytest_gpu.log
tail -10 gpurun_out/r03f_pytest_gpu.log
MVS_RUN_EXPERIMENTS=1 timeout 900 python -m pytest tests/test_gpu_experiments.py -m gpu -x -q > gpurun_out/r03f_pytest_experiments.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r03f_pytest_experiments.log
tail -3 gpurun_out/r03f_pytest_experiments.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r03f_smoke.log 2>&1; tail -2 gpurun_out/r03f_smoke.log
( time timeout 600 python bench.py > gpurun_out/r03f_bench_1gpu.json 2> gpurun_out/r03f_bench_1gpu.err ) 2> gpurun_out/r03f_bench_time.txt; cat gpurun_out/r03f_bench_time.txt
timeout 300 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/r03f_bench_ref.json 2> gpurun_out/r03f_bench_ref.err
