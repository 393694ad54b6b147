#!/bin/bash
# round-2 evidence: bench lines, ncu launch list of the bench command, one --set full capture per top kernel (each after a clean run)
cd /root/repo
python bench.py > gpurun_out/r2_bench_c2_final.json 2> gpurun_out/r2_bench_c2_final.err; echo "bench rc=$?"
python bench.py --workload c3 --only --no-cpu > gpurun_out/r2_bench_c3_1gpu.json 2> gpurun_out/r2_bench_c3_1gpu.err; echo "c3 rc=$?"
python bench.py --impl reference --steps 5 > gpurun_out/r2_bench_c2_reference.json 2> gpurun_out/r2_bench_c2_reference.err; echo "ref rc=$?"
python bench.py --only --no-cpu --steps 2 > gpurun_out/r2_lb.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -k regex:frx_ -c 400 --csv --log-file gpurun_out/r2_launches_final.csv python bench.py --only --no-cpu --steps 2 > gpurun_out/r2_ncu1.log 2>&1
scratch/prof_one.sh r2_enc_final frx_encode_kernel scratch/prof_enc.py 20000
scratch/prof_one.sh r2_dec_final frx_decode_kernel scratch/dec_prof.py
scratch/prof_one.sh r2_enc_c3 frx_encode_kernel scratch/prof_enc3.py 16384 3
rm -f gpurun_out/r2_enc_c3.ncu-rep gpurun_out/r2_dec_final.ncu-rep
ls -la gpurun_out | tail -20
