mkdir -p gpurun_out
timeout 300 python tools/e2e_probe.py > gpurun_out/r02j_e2e_probe.jsonl 2>&1
cat gpurun_out/r02j_e2e_probe.jsonl
