set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "full_config or heavy_tailed" --durations=5 > gpurun_out/r02w_pytest_full_size.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r02w_pytest_full_size.log
tail -15 gpurun_out/r02w_pytest_full_size.log
