"""InstanceNorm++ (+ ELU) kernels (first piece) against the HBM roofline and against the ATen op
sequence they replace, at the activation shapes of configs 4 / 5 (B = 128 -> 256 stacked rows); then the config-4 / config-5
training step with and without them (FDSM_NORM_KERNEL).  CUDA events, 5 warm-up + 20 timed launches, the tensors (134 MB
each at the first shape) larger than L2.   python scripts/bench_norm.py [--steps-only | --kernels-only]"""
import json
import os
import subprocess
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import fdsm_b200 as fdsm  # noqa: E402

dev = torch.device("cuda:0")
peak = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"] if os.path.exists(os.path.join(ROOT, "MEASURED_PEAKS.json")) else 6543.1


def timed(fn, n=20, warm=5):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n * 1e3      # us


def kernels():
    print(f"HBM peak (MEASURED_PEAKS.json): {peak:.0f} GB/s.  algorithmic bytes: forward 8 B/element (read x, write out), "
          f"backward 16 B/element (read x, out, gout; write dx); executed: 12 and 28")
    print(f"{'shape':>22s} {'rows':>7s} | {'fwd us':>8s} {'GB/s':>7s} {'frac':>5s} {'ATen us':>8s} | {'bwd us':>8s} {'GB/s':>7s} {'frac':>5s} {'ATen us':>8s}")
    for shape in ((256, 128, 32, 32), (256, 256, 16, 16), (256, 256, 8, 8), (1024, 128, 32, 32)):
        for per_sample in (False, True):
            N, C, H, W = shape
            x = torch.randn(*shape, device=dev)
            gout = torch.randn(*shape, device=dev)
            rows = N if per_sample else 1
            gamma, alpha, beta = (torch.randn(rows, C, device=dev) * 0.1 + 1 for _ in range(3))
            res = {}
            for name, fn in (("ours", fdsm.ops.inorm_plus), ("aten", fdsm.ops.inorm_plus_reference)):
                xt, g, a, b = (t.clone().requires_grad_() for t in (x, gamma, alpha, beta))
                with torch.no_grad():
                    res[name + "_fwd"] = timed(lambda: fn(xt, g, a, b, True))
                out = fn(xt, g, a, b, True)
                res[name + "_bwd"] = timed(lambda: torch.autograd.grad(out, (xt, g, a, b), gout, retain_graph=True))
                del out, xt
            el = x.numel()
            f, bw = res["ours_fwd"], res["ours_bwd"]
            print(f"{str(shape):>22s} {'(N,C)' if per_sample else '(1,C)':>7s} | {f:8.1f} {8 * el / f / 1e3:7.0f} {8 * el / f / 1e3 / peak:5.2f} "
                  f"{res['aten_fwd']:8.1f} | {bw:8.1f} {16 * el / bw / 1e3:7.0f} {16 * el / bw / 1e3 / peak:5.2f} {res['aten_bwd']:8.1f}", flush=True)
            del x, gout
            torch.cuda.empty_cache()


def steps():
    for cfg in (5, 4):
        for flag in ("0", "1"):
            env = dict(os.environ, FDSM_NORM_KERNEL=flag)
            r = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--config", str(cfg), "--steps", "10", "--warmup", "3", "--no-cpu-baseline"],
                               env=env, capture_output=True, text=True)
            try:
                d = json.loads(r.stdout.strip().splitlines()[-1])
                print(f"config {cfg}, FDSM_NORM_KERNEL={flag}: {d['ms_per_step']:.2f} ms per step, {d['value']:.1f} samples/s, launches of ours {d.get('gpu_launches')}", flush=True)
            except Exception:
                print(f"config {cfg}, FDSM_NORM_KERNEL={flag}: bench failed\n{r.stderr[-2000:]}", flush=True)


if __name__ == "__main__":
    if "--steps-only" not in sys.argv:
        kernels()
    if "--kernels-only" not in sys.argv:
        steps()
