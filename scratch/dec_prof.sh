#!/bin/bash
cd /root/repo
python scratch/dec_prof.py > gpurun_out/fp_dec.log 2>&1 || exit 1
ncu --set full --clock-control none --import-source on -k regex:frx_decode_kernel --launch-skip 2 --launch-count 1 -o /tmp/dec -f python scratch/dec_prof.py > gpurun_out/fp_ncu_dec.log 2>&1
python scratch/ncu_lines.py /tmp/dec.ncu-rep 60 > gpurun_out/r1_dec_final_ncu_summary.txt 2>&1
ncu -i /tmp/dec.ncu-rep --page details > gpurun_out/r1_dec_final_ncu_details.txt 2>&1
python scratch/ncu_regions.py /tmp/dec.ncu-rep /root/repo/flac_raster_b200/csrc/frx_decode_frame.cuh > gpurun_out/dec_regions.txt 2>&1
head -75 gpurun_out/r1_dec_final_ncu_summary.txt
