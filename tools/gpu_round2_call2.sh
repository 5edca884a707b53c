set -x
mkdir -p gpurun_out
python tools/cubin_key.py > gpurun_out/r02b_cubin_key.txt
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/r02b_pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r02b_pytest_gpu.log
tail -3 gpurun_out/r02b_pytest_gpu.log
timeout 600 python bench.py --no-cpu-baseline > gpurun_out/r02b_bench_1gpu.json 2> gpurun_out/r02b_bench_1gpu.err; echo "bench rc=$?"
for d in 256 512 2048; do timeout 300 python tools/run_shape.py --d $d --steps 10; done > gpurun_out/r02b_shapes.jsonl 2>&1
timeout 300 ./tools/kexp/kexp 26000 3840 6 mix > gpurun_out/r02b_kexp.txt 2>&1
