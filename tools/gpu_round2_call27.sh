set -x
mkdir -p gpurun_out
python tools/cubin_key.py > gpurun_out/r03e_cubin_key.txt
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_zz_no_alloc.py -m gpu -x -q > gpurun_out/r03e_pytest_parity.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r03e_pytest_parity.log
tail -4 gpurun_out/r03e_pytest_parity.log
timeout 600 python bench.py --no-cpu-baseline > gpurun_out/r03e_bench_1gpu.json 2> gpurun_out/r03e_bench_1gpu.err; echo "bench rc=$?"
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r03e_launches.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-config3 --c4-samples 0 > gpurun_out/r03e_ncu_launches.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:project_kernel -s 3 -c 1 -f -o gpurun_out/r03e_project_full python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-config3 --c4-samples 0 > gpurun_out/r03e_ncu_full.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:project_kernel -s 3 -c 1 -f -o gpurun_out/r03e_project_d256_full python tools/run_shape.py --d 256 > gpurun_out/r03e_ncu_d256.log 2>&1
ls -la gpurun_out | grep r03a
