#!/bin/bash
# round-end evidence: launch list of the bench command + one --set full capture per top kernel (each after a clean run without ncu);
# the reports are summarised on the box (source-level hot lines + details page) and only the summaries travel back
set -x
cd /root/repo
python bench.py --steps 2 --warmup 3 --no-cpu > gpurun_out/fp_bench.log 2>&1 || exit 1
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r1_launches_final.csv python bench.py --steps 2 --warmup 3 --no-cpu > gpurun_out/fp_ncu1.log 2>&1
cap() {  # cap <name> <kernel regex> <script...>
  name=$1; kern=$2; shift 2
  python "$@" > gpurun_out/fp_$name.log 2>&1 || return 1
  ncu --set full --clock-control none --import-source on -k regex:$kern --launch-skip 2 --launch-count 1 -o /tmp/$name -f python "$@" > gpurun_out/fp_ncu_$name.log 2>&1
  python scratch/ncu_lines.py /tmp/$name.ncu-rep 45 > gpurun_out/r1_${name}_final_ncu_summary.txt 2>&1
  ncu -i /tmp/$name.ncu-rep --page details > gpurun_out/r1_${name}_final_ncu_details.txt 2>&1
}
cap enc frx_encode_kernel scratch/prof_enc.py 20000
cap dec frx_decode_kernel scratch/dec_prof.py
cap ing frx_tiff_unpack_kernel scratch/ing_prof.py
ls -la gpurun_out /tmp/*.ncu-rep
for f in gpurun_out/fp_*.log; do echo "== $f"; tail -n 2 $f; done
