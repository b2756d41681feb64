#!/bin/bash
# Runs bench.py once per environment setting ("VAR=val[,VAR2=val2]") and prints kernel ms (tuning aid, not a bench number).
for v in "$@"; do
  env $(echo "$v" | tr ',' ' ') python bench.py --steps 5 --warmup 3 --no-cpu-baseline 2>/dev/null | tail -1 | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print('$v', 'kernel_ms=%.3f'%d['roofline']['kernel_ms'], 'frac=%.3f'%d['roofline']['frac'], 'step_ms=%.3f'%d['ms_per_step'], d['stages_ms'])"
done
