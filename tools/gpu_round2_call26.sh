set -x
mkdir -p gpurun_out
timeout 600 python tools/ksweep_c3.py > gpurun_out/r02z_ksweep_c3.jsonl 2> gpurun_out/r02z_ksweep_c3.err; echo "rc=$?"; tail -3 gpurun_out/r02z_ksweep_c3.err
