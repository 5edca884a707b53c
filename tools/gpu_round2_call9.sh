mkdir -p gpurun_out
timeout 300 python tools/h2d_align_probe.py > gpurun_out/r02i_h2d_align_probe.json 2>&1
cat gpurun_out/r02i_h2d_align_probe.json
