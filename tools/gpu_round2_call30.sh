set -x
mkdir -p gpurun_out
port=29590
for n in 8 4 2; do
port=$((port+1))
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $port bench.py --gpus $n > gpurun_out/r03h_bench_${n}gpu.json 2> gpurun_out/r03h_bench_${n}gpu.err; echo "bench rc=$?"
done
