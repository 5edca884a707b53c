set -x
mkdir -p gpurun_out
for lib in libmvs_b200_exp.so libmvs_b200_exp_trip2s.so libmvs_b200_exp.so libmvs_b200_exp_trip2s.so; do
MVS_EXP_LIB=$PWD/metagenome_vector_sketches_b200/$lib timeout 300 python tools/ab_c3.py >> gpurun_out/r03g_ab_c3.jsonl 2>> gpurun_out/r03g_ab_c3.err
done
cat gpurun_out/r03g_ab_c3.jsonl
