set -x
mkdir -p gpurun_out
python tools/cubin_key.py > gpurun_out/r02f_cubin_key.txt
timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -x -q > gpurun_out/r02f_pytest_parity.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r02f_pytest_parity.log
tail -3 gpurun_out/r02f_pytest_parity.log
for d in 256 512 1024 2048 4096; do timeout 300 python tools/run_shape.py --d $d --steps 10; done > gpurun_out/r02f_shapes.jsonl 2>&1
cat gpurun_out/r02f_shapes.jsonl
timeout 600 python bench.py --no-cpu-baseline > gpurun_out/r02f_bench_1gpu.json 2> gpurun_out/r02f_bench_1gpu.err; echo "bench rc=$?"
