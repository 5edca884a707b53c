set -x
mkdir -p gpurun_out
timeout 300 ./tools/kexp/kexp 26000 3840 6 > gpurun_out/r02c_kexp.txt 2>&1
cat gpurun_out/r02c_kexp.txt
