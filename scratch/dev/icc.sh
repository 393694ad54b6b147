#!/bin/bash
# scratch/dev/icc.sh <tag>...: instruction-cache and issue metrics of the encode kernel (8192^2 DEM)
for t in "$@"; do
ncu --metrics sm__icc_request_hit_rate.pct,sm__icc_requests.sum,gcc__cache_requests_type_instruction.sum,gcc__cache_requests_type_instruction.sum.pct_of_peak_sustained_elapsed,smsp__issue_active.avg.pct_of_peak_sustained_active,gpu__time_duration.sum,smsp__inst_executed.sum --clock-control none -k regex:frx_encode_kernel --launch-skip 2 --launch-count 1 python scratch/dev/prof.py $t 2>&1 | grep -E "icc|gcc|issue_active|duration|inst_executed" | sed "s/^/$t /"
done
