"""Oracle restatement of the sampler / EMA callers against the reference's own outputs."""
import numpy as np
import torch

from oracle import cases, nets, sampling


def _setup(dt):
    sig = cases.geometric_sigmas(1.0, 0.1, 3)
    net = nets.TinyCondScore((3, 8, 8), 40, 3, sigmas=sig)
    nets.init_from_numpy(net, 4242)
    x0 = nets.np_rand(4243, 6, 3, 8, 8).to(dt)
    zs = [nets.np_randn(4300 + i, 6, 3, 8, 8).to(dt) for i in range(6)]
    return net.to(dt), x0, zs, sig


def test_langevin_restatement_equals_reference(golden):
    for tag, dt in (("f32", torch.float32), ("f64", torch.float64)):
        net, x0, zs, sig = _setup(dt)
        x = sampling.anneal_langevin(x0, net, sig.numpy(), 2, 2e-3, zs)
        np.testing.assert_array_equal(x.double().numpy(), golden[f"n3_langevin/x_final_{tag}"])


def test_ema_restatement_equals_reference(golden):
    for tag, dt in (("f32", torch.float32), ("f64", torch.float64)):
        lin = torch.nn.Linear(7, 5).to(dt)
        nets.init_from_numpy(lin, 4400)
        shadow = [p.data.clone() for p in lin.parameters()]
        for i in range(3):
            with torch.no_grad():
                for j, p in enumerate(lin.parameters()):
                    p.add_(nets.np_randn(4410 + 10 * i + j, *p.shape).to(dt) * 0.1)
            shadow = [sampling.ema_update(s, p.data, 0.9) for s, p in zip(shadow, lin.parameters())]
        np.testing.assert_array_equal(torch.cat([s.reshape(-1) for s in shadow]).double().numpy(), golden[f"n4_ema/shadow_{tag}"])


def test_oracle_mdsm_sampler_matches_reference_golden():
    """oracle.sampling.mdsm_annealed_langevin against outputs of the unmodified DeepEBM/ESM/sampling_mdsm.py
    (tests/golden/deepebm_sampler.npz, oracle/make_golden_sampler.py)."""
    import os
    import numpy as np
    import torch
    from oracle import sampling as osamp
    from oracle.make_golden_sampler import energy_net, inputs
    g = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "deepebm_sampler.npz"))
    for tag, dt in (("f32", torch.float32), ("f64", torch.float64)):
        net = energy_net(dt)
        x0, zs, T = inputs(dt)
        x, E = osamp.mdsm_annealed_langevin(net, x0.clone(), 0.1, T, zs)
        np.testing.assert_array_equal(x.double().numpy(), g[f"annealed/x_final_{tag}"])
        np.testing.assert_array_equal(E.double().numpy(), g[f"annealed/E_{tag}"])
        x, _ = osamp.mdsm_annealed_langevin(net, x0.clone(), 0.1, np.ones(4), zs[:4])
        np.testing.assert_array_equal(x.double().numpy(), g[f"langevin/x_final_{tag}"])
