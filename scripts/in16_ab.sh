#!/bin/bash
# Config-2 step with the input layer in TF32 (jet_passes 2) and in kind::f16 (jet_passes 4).  bash scripts/in16_ab.sh
run() { env "$@" timeout 300 python bench.py --no-cpu-baseline 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('$*', '| step %.1f us  chain %.1f us  u8 %.1f' % (d['ms_per_step']*1e3, d['roofline']['avg_launch_us'], d['e2e_u8']['ms_per_step']*1e3))"; }
run FDSM_BENCH_JET_PASSES=2
run FDSM_BENCH_JET_PASSES=4
