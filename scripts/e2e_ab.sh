#!/bin/bash
# A/B of the end-to-end step's host plumbing (train.py step_from_host): loss read by kernel store vs memcpy D2H, one vs two
# H2D copy streams.  bash scripts/e2e_ab.sh  -> gpurun_out/e2e_ab.txt
out=gpurun_out/e2e_ab.txt; : > $out
run() {
  env "$@" python bench.py --steps 200 --warmup 20 --no-cpu-baseline 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1])
print('%-60s step %.1f us | e2e fp32 %.1f us (%.2f M/s) | e2e u8 %.1f us (%.2f M/s) | h2d %.1f GB/s' % ('$*', d['ms_per_step']*1e3, d['e2e']['ms_per_step']*1e3, d['e2e']['value']/1e6, d['e2e_u8']['ms_per_step']*1e3, d['e2e_u8']['value']/1e6, d['e2e']['h2d_gbs_overlapped']))" >> $out
}
run FDSM_E2E_LOSS_MEMCPY=1 FDSM_E2E_COPY_STREAMS=1
run FDSM_E2E_LOSS_MEMCPY=0 FDSM_E2E_COPY_STREAMS=2
run FDSM_E2E_LOSS_MEMCPY=1 FDSM_E2E_COPY_STREAMS=2
run FDSM_E2E_LOSS_MEMCPY=0 FDSM_E2E_COPY_STREAMS=1
cat $out
