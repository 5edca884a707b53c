set -x
mkdir -p gpurun_out
nvidia-smi -L > gpurun_out/r02a_gpu.txt; lscpu | head -20 >> gpurun_out/r02a_gpu.txt
python tools/cubin_key.py > gpurun_out/r02a_cubin_key.txt
timeout 900 python -m pytest tests -m gpu -x -q --durations=15 > gpurun_out/r02a_pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r02a_pytest_gpu.log
MVS_RUN_EXPERIMENTS=1 timeout 900 python -m pytest tests/test_gpu_experiments.py -m gpu -x -q > gpurun_out/r02a_pytest_experiments.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r02a_pytest_experiments.log
timeout 600 python bench.py > gpurun_out/r02a_bench_1gpu.json 2> gpurun_out/r02a_bench_1gpu.err; echo "bench rc=$?"
timeout 300 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/r02a_bench_ref.json 2> gpurun_out/r02a_bench_ref.err
timeout 300 ./tools/kexp/kexp 26000 3840 6 > gpurun_out/r02a_kexp.txt 2>&1
timeout 300 python tools/run_shape.py --d 256 > gpurun_out/r02a_shape_d256.json 2>&1
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r02a_launches.csv python bench.py --steps 2 --warmup 1 --no-cpu-baseline --no-config3 --c4-samples 0 > gpurun_out/r02a_ncu_launches.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:project_kernel -s 3 -c 1 -f -o gpurun_out/r02a_project_full python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-config3 --c4-samples 0 > gpurun_out/r02a_ncu_full.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:project_kernel -s 3 -c 1 -f -o gpurun_out/r02a_project_d256_full python tools/run_shape.py --d 256 > gpurun_out/r02a_ncu_d256.log 2>&1
ls -la gpurun_out | tail -20
