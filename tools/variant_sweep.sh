#!/bin/bash
# Runs bench.py for each QP kernel variant and prints kernel ms (tuning aid, not a bench number).
for v in "$@"; do
  MRB_QP_VARIANT=$v python bench.py --steps 3 --warmup 2 --no-cpu-baseline 2>/dev/null | tail -1 | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print('$v', 'kernel_ms=%.3f'%d['roofline']['kernel_ms'], 'frac=%.3f'%d['roofline']['frac'], 'step_ms=%.3f'%d['ms_per_step'], d['stages_ms'])"
done
