set -x
mkdir -p gpurun_out
# CLI tests after the in-tree shard writers and the binary hash-file input; CLI wall clock from text and from the binary form
timeout 900 python -m pytest tests/test_gpu_cli.py -m gpu -x -q --durations=5 > gpurun_out/r02u_pytest_gpu_cli.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r02u_pytest_gpu_cli.log
tail -12 gpurun_out/r02u_pytest_gpu_cli.log
timeout 900 python tests/perf_cli.py > gpurun_out/r02u_cli_bench.json 2> gpurun_out/r02u_cli_bench.err; echo "perf rc=$?"
tail -c 1500 gpurun_out/r02u_cli_bench.json
