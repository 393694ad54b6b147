#!/bin/bash
cd /root/repo
python -m pytest tests -m gpu -x -q 2>&1 | tail -3
python bench.py > gpurun_out/r1_bench_c2_s11.json 2> gpurun_out/s11_b.err
python bench.py --impl reference > gpurun_out/r1_bench_c2_reference_s11.json 2>> gpurun_out/s11_b.err
python bench.py --workload c3 --no-cpu > gpurun_out/r1_bench_c3_s11.json 2>> gpurun_out/s11_b.err
tail -3 gpurun_out/s11_b.err
for f in gpurun_out/r1_bench_c2_s11.json gpurun_out/r1_bench_c3_s11.json; do python -c "
import json,sys; d=json.load(open('$f')); print(d['value'], d['ms_per_step'], d['roofline']['kernel_ms'], d['e2e'], d['decode']['value'], d['decode']['roofline']['kernel_ms'], d['ingest']['value'], d.get('cpu_baseline'))"; done
cat gpurun_out/r1_bench_c2_reference_s11.json
