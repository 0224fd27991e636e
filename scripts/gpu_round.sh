#!/bin/bash
# One GPU-box pass of the round's evidence: full GPU test log, N=1 bench (driver-style and long), reference arm,
# ncu launch list of the bench command and one `ncu --set full` capture of the dominant kernel.
#   gpurun --timeout 1500 -- 'bash scripts/gpu_round.sh r02'
tag=${1:-r02}
out=gpurun_out
mkdir -p $out
nvidia-smi --query-gpu=name,clocks.max.sm,clocks.sm,power.limit --format=csv > $out/${tag}_gpu.txt 2>&1
timeout 900 python -m pytest tests -m gpu -x -q > $out/${tag}_pytest_gpu.txt 2>&1; echo "pytest rc=$?" | tee -a $out/${tag}_pytest_gpu.txt
tail -3 $out/${tag}_pytest_gpu.txt
timeout 600 python bench.py --steps 20 --warmup 5 > $out/${tag}_bench_n1_driver_args.json 2> $out/${tag}_bench_n1_driver_args.err; echo "bench rc=$?"
timeout 600 python bench.py --no-cpu-baseline > $out/${tag}_bench_n1.json 2> $out/${tag}_bench_n1.err; echo "bench(200) rc=$?"
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > $out/${tag}_bench_reference_arm.json 2> $out/${tag}_bench_reference_arm.err; echo "ref rc=$?"
timeout 600 python -c "import __graft_entry__ as g; g.smoke()" > $out/${tag}_smoke.txt 2>&1; echo "smoke rc=$?"
timeout 300 python scripts/chain_trace.py 4096 4 > $out/${tag}_chain_trace.txt 2>&1; echo "chain trace rc=$?"
timeout 300 python scripts/update_trace.py > $out/${tag}_update_trace.txt 2>&1; echo "update trace rc=$?"
timeout 300 python scripts/bench_fd_vs_exact.py > $out/${tag}_fd_vs_exact.txt 2>&1; echo "fd vs exact rc=$?"
for c in 1 3 4 5; do timeout 600 python bench.py --config $c --no-cpu-baseline > $out/${tag}_bench_cfg$c.json 2> $out/${tag}_bench_cfg$c.err; echo "cfg$c rc=$?"; done
timeout 300 python scripts/bench_norm.py --kernels-only > $out/${tag}_norm_kernels.txt 2>&1; echo "norm kernels rc=$?"
export FDSM_BENCH_SUSTAIN=0
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file $out/${tag}_ncu_launches_bench.csv \
  python bench.py --steps 2 --warmup 3 --no-cpu-baseline > $out/${tag}_ncu_launches.log 2>&1; echo "ncu list rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'chain_kernel|jet_|gemm' -c 24 -f -o $out/${tag}_chain_full \
  python bench.py --steps 2 --warmup 3 --no-cpu-baseline > $out/${tag}_ncu_full.log 2>&1; echo "ncu full rc=$?"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:'inorm_(stats|apply)' -c 4 -f -o $out/${tag}_norm_fwd_full \
  python scripts/bench_norm.py --kernels-only > $out/${tag}_ncu_norm.log 2>&1; echo "ncu norm fwd rc=$?"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:'inorm_bwd' -c 4 -f -o $out/${tag}_norm_bwd_full \
  python scripts/bench_norm.py --kernels-only >> $out/${tag}_ncu_norm.log 2>&1; echo "ncu norm bwd rc=$?"
ls -la $out
