#!/bin/bash
# prof_one.sh <name> <kernel regex> <script> [args]: clean run, then one --set full capture, summaries into gpurun_out/
cd /root/repo
name=$1; kern=$2; shift 2
python "$@" > gpurun_out/${name}_run.log 2>&1 || { tail -5 gpurun_out/${name}_run.log; exit 1; }
ncu --set full --clock-control none --import-source on -k regex:$kern --launch-skip 2 --launch-count 1 -o /tmp/$name -f python "$@" > gpurun_out/${name}_ncu.log 2>&1
python scratch/ncu_lines.py /tmp/$name.ncu-rep 60 > gpurun_out/${name}_ncu_summary.txt 2>&1
python scratch/ncu_spills.py /tmp/$name.ncu-rep > gpurun_out/${name}_ncu_spills.txt 2>&1
ncu -i /tmp/$name.ncu-rep --page details > gpurun_out/${name}_ncu_details.txt 2>&1
cp /tmp/$name.ncu-rep gpurun_out/ 2>/dev/null; ls -la gpurun_out/$name.ncu-rep; cat gpurun_out/${name}_run.log
