#!/bin/bash
# Critical-path share of each kernel of the fused step: the step is timed with one kernel left out (FDSM_JET_SKIP bit mask;
# the results of such a run are garbage, only the time is read).  bash scripts/step_shares.sh
for m in 0 1 2 4 8 14 7; do
  FDSM_JET_SKIP=$m python bench.py --steps 50 --warmup 10 --no-cpu-baseline 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('skip mask $m: step %.1f us' % (d['ms_per_step']*1e3))"
done
