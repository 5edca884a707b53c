set -x
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "carry_window or random_sets" > gpurun_out/r03i_pytest_carry.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r03i_pytest_carry.log
tail -5 gpurun_out/r03i_pytest_carry.log
