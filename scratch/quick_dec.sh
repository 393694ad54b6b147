#!/bin/bash
cd /root/repo
python -m pytest tests -m gpu -x -q 2>&1 | tail -2
python bench.py --no-cpu 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print('c2', d['value'], d['ms_per_step'], d['decode']['value'], d['decode']['roofline']['kernel_ms'])"
python bench.py --workload c3 --no-cpu 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print('c3', d['value'], d['ms_per_step'], d['decode']['value'], d['decode']['roofline']['kernel_ms'])"
python scratch/c5_bench.py 2>&1 | tail -1
