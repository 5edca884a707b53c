mkdir -p gpurun_out
P=metagenome_vector_sketches_b200
for lib in libmvs_b200_exp_old.so libmvs_b200_exp_trip1.so libmvs_b200_exp_trip2.so libmvs_b200_exp.so; do
  MVS_EXP_LIB=$PWD/$P/$lib timeout 300 python tools/ab_kernel.py 2>&1 | grep -v Warning
done > gpurun_out/r02e_ab_kernel.jsonl
cat gpurun_out/r02e_ab_kernel.jsonl
