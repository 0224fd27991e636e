#!/bin/bash
# Multi-GPU step, A/B of the exchange variants.  bash scripts/ddp_ab.sh N  -> gpurun_out/ddp_ab_nN.txt
N=${1:-2}
out=gpurun_out/ddp_ab_n$N.txt; : > $out
run() {
  env "$@" python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29741 bench.py --gpus $N --steps 200 --warmup 20 --no-cpu-baseline 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1])
c=d.get('ddp_check') or {}
print('%-50s N=%d step %.1f us  %.2f M samples/s  exchange %s us  e2e %.1f us  ddp_check all ok: %s' % ('$*', d['n_gpus'], d['ms_per_step']*1e3, d['value']/1e6, d.get('exchange_us'), d['e2e']['ms_per_step']*1e3, c.get('all_ranks_ok')))
open('gpurun_out/ddp_ab_n$N.' + '$*'.replace(' ', '_').replace('=', '') + '.json', 'w').write(json.dumps(d))" >> $out
}
run FDSM_ALLREDUCE=peer
run FDSM_ALLREDUCE=nccl
[ "$N" -le 4 ] && run FDSM_PEER_TWO_SHOT=1
cat $out
