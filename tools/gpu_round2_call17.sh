set -x
mkdir -p gpurun_out
# the allocation-free entry point: alone in a fresh process, then after the whole GPU suite
tail -5 gpurun_out/r02r_pytest_no_alloc_alone.log
timeout 900 python -m pytest tests -m gpu -x -q --durations=6 > gpurun_out/r02r_pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r02r_pytest_gpu.log
tail -12 gpurun_out/r02r_pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r02r_smoke.log 2>&1; tail -2 gpurun_out/r02r_smoke.log
