#!/bin/bash
# round-end evidence: launch list of the bench command + one --set full capture per top kernel (each after a clean run without ncu)
set -x
cd /root/repo
python bench.py --steps 2 --warmup 3 --no-cpu > gpurun_out/fp_bench.log 2>&1 || exit 1
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r1_launches_final.csv python bench.py --steps 2 --warmup 3 --no-cpu > gpurun_out/fp_ncu1.log 2>&1
python scratch/prof_enc.py 20000 > gpurun_out/fp_enc.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:frx_encode_kernel --launch-skip 2 --launch-count 1 -o gpurun_out/r1_enc_final -f python scratch/prof_enc.py 20000 > gpurun_out/fp_ncu2.log 2>&1
python scratch/dec_prof.py > gpurun_out/fp_dec.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:frx_decode_kernel --launch-skip 2 --launch-count 1 -o gpurun_out/r1_dec_final -f python scratch/dec_prof.py > gpurun_out/fp_ncu3.log 2>&1
python scratch/ing_prof.py > gpurun_out/fp_ing.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:frx_tiff_unpack_kernel --launch-skip 2 --launch-count 1 -o gpurun_out/r1_ing_final -f python scratch/ing_prof.py > gpurun_out/fp_ncu4.log 2>&1
tail -2 gpurun_out/fp_*.log
