"""bench.py --config {1,3,4,5}: the other BASELINE.json configurations (config 2 is bench.py's default line).

    1  FD-SSM (efficient_score_matching_conjugate, eps 0.1) on the Softplus MLP 11-128-128-1, GMM data, B=1024
       -- fused mode (tcgen05 MLP); the reference's CPU-runnable case, reported on the GPU as well
    3  DeepEBM ssm_fd on Res18_Quadratic(1, 128, 32), synthetic 28x28 images zero-padded to 32x32, B=64/GPU
    4  ncsn anneal_ESM_scorenet_VR (10 sigmas 1 -> 0.01) on CondRefineNetDilated (ngf 128), x (128, 3, 32, 32)
    5  ncsnv2 anneal_ESM_scorenet_VR (232 sigmas 50 -> 0.01, anneal_power 2) on NCSNv2 (ngf 128), B=128/GPU, EMA 0.999

Configs 3-5 run in GENERIC mode: kernels (1)+(2) and (4) of libfdsm around the torch network (cuDNN; SURVEY.md section
8d: "library-backed; reported, not judged").  The JSON line reports, next to samples/s, the share of the step our FD
kernels take and their HBM roofline at the config's shape.
"""
import math
import os
import time

import torch

CONFIGS = {
    1: dict(name="cfg1: SSM efficient_score_matching_conjugate, Softplus MLP 11-128-128-1, GMM(11), B=1024/GPU, Adam 1e-3",
            B=1024, shape=(11,), steps=200, warmup=20),
    3: dict(name="cfg3: DeepEBM ssm_fd, Res18_Quadratic(1,128,32) elu, 28x28 padded to 32x32, B=64/GPU, Adam 5e-5 + clip 0.1",
            B=64, shape=(1, 32, 32), steps=10, warmup=3),
    4: dict(name="cfg4: ncsn anneal_ESM_scorenet_VR (10 sigmas), CondRefineNetDilated ngf128, B=128/GPU x (3,32,32), Adam 1e-3",
            B=128, shape=(3, 32, 32), steps=10, warmup=3),
    5: dict(name="cfg5: ncsnv2 anneal_ESM_scorenet_VR (232 sigmas, anneal_power 2), NCSNv2 ngf128, B=128/GPU x (3,32,32), Adam 1e-4 + EMA",
            B=128, shape=(3, 32, 32), steps=10, warmup=3),
}


def synth(cfg, gen, device, n=None):
    B = cfg["B"] if n is None else n
    if cfg["id"] == 1:      # GMMDist(11): means +-5, weights .8/.2, sigma 1 (ncsn/models/gmm.py:58-73)
        sign = (torch.rand(B, 1, generator=gen, device=device) < 0.8).float() * 2 - 1
        return torch.randn(B, 11, generator=gen, device=device) + 5.0 * sign
    if cfg["id"] == 3:      # MNIST-range pixels, 28x28 zero-padded to 32x32 (DeepEBM/ESM/data.py:156) + N(0, 0.005^2)
        x = torch.zeros(B, 1, 32, 32, device=device)
        x[:, :, 2:30, 2:30] = torch.rand(B, 1, 28, 28, generator=gen, device=device)
        return x + 0.005 * torch.randn(B, 1, 32, 32, generator=gen, device=device)
    # CIFAR-range pixels, uniformly dequantised (ncsn/runners/anneal_runner.py:137)
    x = torch.randint(0, 256, (B, 3, 32, 32), generator=gen, device=device).float()
    return (x + torch.rand(B, 3, 32, 32, generator=gen, device=device)) / 256.0


def build(cfg_id, device, fdsm):
    """-> (cfg, net, trainer)"""
    from fdsm_b200 import losses, nets
    from fdsm_b200.train import FDSSMTrainer, GenericFDTrainer
    cfg = dict(CONFIGS[cfg_id], id=cfg_id)
    torch.manual_seed(1234)
    if cfg_id in (3, 4, 5) and os.environ.get("FDSM_BENCH_CUDNN_BENCHMARK", "1") != "0":
        # as the reference's runners run (DeepEBM/train_ssm.py:23, ncsn/main.py:96, ncsnv2/main.py:193)
        torch.backends.cudnn.benchmark = True
    if cfg_id == 1:
        net = fdsm.mlp.FusedEnergyMLP(11, 128, 2, act="softplus", head="linear", device=device, seed=1234)
        tr = FDSSMTrainer(net, eps=0.1, lr=1e-3, betas=(0.9, 0.999), clip=None, seed=1234, scale=11 / 0.1 ** 2)
        cfg["network"] = "fused tcgen05 MLP (ours), one CUDA-graph replay per step"
    elif cfg_id == 3:
        net = nets.Res18Quadratic(1, 128, 32, act="elu").to(device)
        tr = GenericFDTrainer(net, lambda m, x: losses.ssm_fd(m, x, eps=0.05, batch_size=x.shape[0]), lr=5e-5, betas=(0.9, 0.95),
                              clip=0.1)
        cfg["network"] = "torch Res18Quadratic (cuDNN) between our FD kernels"
    elif cfg_id == 4:
        net = nets.CondRefineNetDilated(3, 128, 10).to(device)
        sig = nets.geometric_sigmas(1.0, 0.01, 10).to(device)

        def loss4(m, x):
            labels = torch.randint(0, 10, (x.shape[0],), device=x.device)       # anneal_runner.py:144
            return losses.ncsn_anneal_ESM_scorenet_VR(m, x, labels, sig, 1, 0.1)
        tr = GenericFDTrainer(net, loss4, lr=1e-3, betas=(0.9, 0.999))
        cfg["network"] = "torch CondRefineNetDilated: cuDNN convolutions, our InstanceNorm++ / ELU kernels, between our FD kernels"
    elif cfg_id == 5:
        net = nets.NCSNv2(3, 128, 232).to(device)
        tr = GenericFDTrainer(net, lambda m, x: losses.anneal_ESM_scorenet_VR(m, x, m.sigmas, None, 2.0, None, 1, 0.1),
                              lr=1e-4, betas=(0.9, 0.999), ema_rate=0.999)
        cfg["network"] = "torch NCSNv2: cuDNN convolutions, our InstanceNorm++ / ELU kernels, between our FD kernels"
    else:
        raise SystemExit(f"unknown --config {cfg_id}")
    return cfg, net, tr


def fd_kernel_times(cfg, device, pk):
    """Our FD kernels alone at the config's shape, graph-timed on rotating buffers (no host launch gaps): the stack
    build (kernels (1)+(2)) and, for the score form, the fused combine (kernel (4) fwd+bwd).  Returns a dict for the
    JSON line: per-launch us, algorithmic GB/s (SURVEY.md section 8d figures) and fraction of the measured HBM peak."""
    from fdsm_b200 import ops
    B, D = cfg["B"], int(math.prod(cfg["shape"]))
    score = cfg["id"] in (4, 5)
    R = 8
    xs = [torch.rand(B, D, device=device) for _ in range(R)]
    rows = 2 if score else 3
    outs = [torch.empty(rows * B, D, device=device) for _ in range(R)]
    sig = torch.linspace(1.0, 0.01, 16, device=device)
    labels = torch.randint(0, 16, (B,), device=device)
    vs = [torch.randn(B, D, device=device) * 0.01 for _ in range(R)]
    ss = [torch.randn(2 * B, D, device=device) for _ in range(R)]

    def stack_once(i):
        if score:
            ops.stack(xs[i % R], row_mask=6, scale=0.1, sigma=sig, labels=labels, perturb=True, out=outs[i % R], draw=(7, i, 0))
        else:
            ops.stack(xs[i % R], row_mask=7, scale=0.05, out=outs[i % R], draw=(7, i, 0))

    def combine_once(i):
        ops.combine_score_fused(ss[i % R], vs[i % R], k=D / 0.02, sigma=sig, labels=labels)

    def graph_time(fn, n=20, reps=5):
        gs = torch.cuda.Stream()
        gs.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(gs):
            fn(0)
            g = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g, stream=gs):
                for i in range(n):
                    fn(i)
            g.replay()
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(gs)
            for _ in range(reps):
                g.replay()
            e1.record(gs)
            torch.cuda.synchronize()
        torch.cuda.current_stream().wait_stream(gs)
        return e0.elapsed_time(e1) * 1e3 / (n * reps)

    out = {}
    us = graph_time(stack_once)
    alg = (12.0 if score else 16.0) * D * B
    out["stack"] = {"us": us, "algorithmic_bytes": alg, "gbs": alg / us / 1e3, "frac_hbm": alg / us / 1e3 / pk["hbm_gbs"]}
    total = us
    if score:
        us = graph_time(combine_once)
        alg = 20.0 * D * B
        out["combine_score_fused"] = {"us": us, "algorithmic_bytes": alg, "gbs": alg / us / 1e3,
                                      "frac_hbm": alg / us / 1e3 / pk["hbm_gbs"], "launches": 2}
        total += us
    out["total_us"] = total
    return out


def cpu_baseline(cfg_id, budget_s=15.0):
    """The reference's CPU path (oracle loss restatement + the same torch network on the host cores) on a bounded
    sample: a reduced batch for the conv nets, scaled per sample."""
    from oracle import fd_losses, nets as onets
    from fdsm_b200 import nets
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    cfg = dict(CONFIGS[cfg_id], id=cfg_id)
    b = {1: 1024, 3: 8, 4: 4, 5: 4}[cfg_id]
    g = torch.Generator().manual_seed(1234)
    x = synth(cfg, g, "cpu", n=b)
    if cfg_id == 1:
        net = onets.init_from_numpy(onets.SoftplusMLP(11, 128), 1234)
        lossf = lambda: fd_losses.fd_ssm(net, x, fd_losses.sphere_direction(torch.randn_like(x), 0.1), 11 / 0.01)
        lr = 1e-3
    elif cfg_id == 3:
        net = nets.Res18Quadratic(1, 128, 32)
        lossf = lambda: fd_losses.fd_ssm(net, x.reshape(b, -1), fd_losses.sphere_direction(torch.randn(b, 1024), 0.05), 400.0)
        lr = 5e-5
    else:
        L = 10 if cfg_id == 4 else 232
        net = nets.CondRefineNetDilated(3, 128, 10) if cfg_id == 4 else nets.NCSNv2(3, 128, 232)
        sig = nets.geometric_sigmas(1.0 if cfg_id == 4 else 50.0, 0.01, L)

        def lossf():
            labels = torch.randint(0, L, (b,))
            v = fd_losses.sphere_direction(torch.randn_like(x), 0.1)
            return fd_losses.anneal_fd_ssmvr(net, x, torch.randn_like(x), v, labels, sig, 0.1, 3072.0, 2.0)
        lr = 1e-3 if cfg_id == 4 else 1e-4
    opt = torch.optim.Adam(net.parameters(), lr=lr)

    def step():
        loss = lossf()
        opt.zero_grad()
        loss.backward()
        opt.step()
    step()
    t0, n = time.perf_counter(), 0
    while True:
        step()
        n += 1
        dt = time.perf_counter() - t0
        if dt > budget_s or n >= 100:
            break
    return {"value": b * n / dt, "unit": "samples/s", "cores": cores, "kind": "port",
            "sample": f"{n} steps at B={b} (of the config's B={cfg['B']}) on the torch-CPU oracle port, {dt:.1f}s, {cores} threads"}


def fd_kernel_rooflines(device, pk, shapes=((4096, 784, "cfg2"), (65536, 784, "saturating, 205 MB in"))):
    """HBM rooflines of the three FD kernels of the generic mode -- the 3-row energy stack, the noised 2-row stack of the
    score form and the fused score combine (forward + ds) -- graph-timed (20 launches per replay, no host gaps) on
    rotating inputs AND outputs larger than L2.  Algorithmic bytes per sample: 16 D, 12 D, 20 D."""
    from fdsm_b200 import ops
    out = {}
    for B, D, tag in shapes:
        nbuf = max(3, min(8, int(300e6 // (B * D * 4)) + 1))
        xs = [torch.rand(B, D, device=device) for _ in range(nbuf)]
        o3 = [torch.empty(3 * B, D, device=device) for _ in range(max(2, nbuf // 2))]
        o2 = [torch.empty(2 * B, D, device=device) for _ in range(max(2, nbuf // 2))]
        sig = torch.exp(torch.linspace(3.9, -4.6, 232)).to(device)
        labels = torch.randint(0, 232, (B,), device=device)
        ss = [torch.randn(2 * B, D, device=device) for _ in range(max(2, nbuf // 2))]
        v = torch.randn(B, D, device=device) * 0.01

        def timed(fn, n=20, reps=3):
            gs = torch.cuda.Stream()
            gs.wait_stream(torch.cuda.current_stream())
            with torch.cuda.stream(gs):
                fn(0)
                g = torch.cuda.CUDAGraph()
                with torch.cuda.graph(g, stream=gs):
                    for i in range(n):
                        fn(i)
                g.replay()
                torch.cuda.synchronize()
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record(gs)
                for _ in range(reps):
                    g.replay()
                e1.record(gs)
                torch.cuda.synchronize()
            torch.cuda.current_stream().wait_stream(gs)
            return e0.elapsed_time(e1) * 1e3 / (n * reps)

        kernels = {
            "stack3_energy": (16.0, lambda i: ops.stack(xs[i % nbuf], row_mask=7, scale=0.05, out=o3[i % len(o3)], draw=(7, i, 0))),
            "stack2_noised": (12.0, lambda i: ops.stack(xs[i % nbuf], row_mask=6, scale=0.1, sigma=sig, labels=labels, perturb=True,
                                                        out=o2[i % len(o2)], draw=(7, i, 0))),
            "combine_score_fused": (20.0, lambda i: ops.combine_score_fused(ss[i % len(ss)], v, k=D / 0.02, sigma=sig, labels=labels)),
        }
        for name, (bps, fn) in kernels.items():
            us = timed(fn)
            alg = bps * D * B
            out[f"{name} B={B} D={D} ({tag})"] = {"bound": "hbm", "avg_launch_us": us, "algorithmic_bytes_per_launch": alg,
                                                 "achieved": alg / us / 1e3, "peak": pk["hbm_gbs"], "unit": "GB/s",
                                                 "frac": alg / us / 1e3 / pk["hbm_gbs"]}
        del xs, o3, o2, ss
        torch.cuda.empty_cache()
    return out
