#!/bin/bash
cd /root/repo
python -m pytest tests -m gpu -x -q 2>&1 | tail -2
python bench.py --no-cpu > gpurun_out/b3.json 2>gpurun_out/b3.err; tail -2 gpurun_out/b3.err
python -c "
import json; d=json.load(open('gpurun_out/b3.json')); print(d['value'], d['ms_per_step'], d['roofline']['kernel_ms'], d['e2e'], d['decode']['value'], d['decode']['ms_per_step'], d['decode']['roofline']['kernel_ms'])"
