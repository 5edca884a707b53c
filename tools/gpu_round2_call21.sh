set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_cli.py -m gpu -x -q > gpurun_out/r02v_pytest_gpu_cli.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r02v_pytest_gpu_cli.log
tail -4 gpurun_out/r02v_pytest_gpu_cli.log
timeout 300 python tools/startup_probe.py > gpurun_out/r02v_startup_probe.json 2> gpurun_out/r02v_startup_probe.err; cat gpurun_out/r02v_startup_probe.json
nvidia-smi --query-gpu=persistence_mode --format=csv
