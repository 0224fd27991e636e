timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/r02g_pytest_gpu.txt 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/r02g_pytest_gpu.txt; tail -3 gpurun_out/r02g_pytest_gpu.txt
timeout 600 python bench.py --steps 20 --warmup 5 > gpurun_out/r02g_bench_n1_driver_args.json 2> gpurun_out/r02g_bench_n1_driver_args.err; echo "bench rc=$?"
timeout 600 python bench.py --no-cpu-baseline > gpurun_out/r02g_bench_n1.json 2> gpurun_out/r02g_bench_n1.err; echo "bench200 rc=$?"
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/r02g_bench_reference_arm.json 2>/dev/null; echo "ref rc=$?"
timeout 600 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r02g_smoke.txt 2>&1; echo "smoke rc=$?"
python - <<'P'
import json
for f in ("r02g_bench_n1_driver_args.json","r02g_bench_n1.json"):
    d=json.loads(open("gpurun_out/"+f).read().strip().splitlines()[-1])
    print(f, "%.1f us %.2f M/s | e2e %.1f us %.2f M/s | u8 %.1f us | chain %.1f us frac %.3f | launches %s" % (d["ms_per_step"]*1e3, d["value"]/1e6, d["e2e"]["ms_per_step"]*1e3, d["e2e"]["value"]/1e6, d["e2e_u8"]["ms_per_step"]*1e3, d["roofline"]["avg_launch_us"], d["roofline"]["frac"], d["gpu_launches"]))
P
