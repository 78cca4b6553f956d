#!/bin/bash
# tests + parity summary + short bench digest (GPU box)
timeout 800 python -m pytest tests -m gpu -q --tb=line 2>&1 | tail -4
python tools/parity_report.py 2>&1 | grep "fp32. WORST"
python bench.py --steps 100 --warmup 10 --no-cpu-baseline --no-e2e "$@" 2>&1 | python -c "
import sys,json
d=json.loads(sys.stdin.read().strip().splitlines()[-1])
print('4096:', round(d['value']/1e6,1),'M env-steps/s', round(d['ms_per_step']*1e3,2),'us', d['config']['launch']['envs_per_block'], d['config']['launch']['threads_per_block'], 'frac', round(d['roofline']['frac'],4))
print([(s['envs'], round(s['ms_per_step']*1e3,1), round(s['hbm_frac'],3), s['launch']) for s in d.get('sweep',[])])"
