"""GPU parity of the Langevin-update and EMA kernels: bit-exact against torch's expression on
injected noise, against the reference's golden outputs through the reference-signature callers, and the in-kernel Philox
mode against the two-step path."""
import numpy as np
import pytest
import torch

from oracle import cases, nets

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


@pytest.mark.parametrize("shape", [(0, 8), (1, 1), (13, 11), (64, 784), (6, 3, 8, 8), (128, 3, 32, 32), (3, 4100)])
def test_langevin_step_bit_exact_with_injected_noise(fdsm, shape):
    x, g, z = (torch.randn(*shape, device=DEV) for _ in range(3))
    step = 3.1e-5
    want = x + step * g + z * float(np.sqrt(step * 2))
    got = fdsm.ops.langevin_step(x, g, step, noise=z, noise_scale=float(np.sqrt(step * 2)))
    assert torch.equal(got, want)
    y = x.clone()
    fdsm.ops.langevin_step(y, g, step, noise=z, noise_scale=float(np.sqrt(step * 2)), out=y)     # in place
    assert torch.equal(y, want)


def test_langevin_philox_mode_equals_two_step(fdsm):
    x, g = torch.randn(256, 3072, device=DEV), torch.randn(256, 3072, device=DEV)
    draw = (77, 5, 1000)
    got = fdsm.ops.langevin_step(x, g, 1e-4, draw=draw)
    z = fdsm.ops.philox_normal(256, 3072, 1, *draw)
    assert torch.equal(got, fdsm.ops.langevin_step(x, g, 1e-4, noise=z))
    assert abs(float(z.mean())) < 5e-3 and abs(float(z.std()) - 1) < 5e-3


def test_anneal_langevin_matches_reference(fdsm, golden):
    sig = cases.geometric_sigmas(1.0, 0.1, 3)
    net = nets.TinyCondScore((3, 8, 8), 40, 3, sigmas=sig)
    nets.init_from_numpy(net, 4242)
    net = net.to(DEV)
    net.sigmas = net.sigmas.to(DEV)
    x0 = nets.np_rand(4243, 6, 3, 8, 8).float().to(DEV)
    zs = [nets.np_randn(4300 + i, 6, 3, 8, 8).float().to(DEV) for i in range(6)]
    out = fdsm.sampling.anneal_Langevin_dynamics(x0, net, sig.numpy(), n_steps_each=2, step_lr=2e-3, final_only=True, noises=zs)
    assert len(out) == 1 and out[0].device.type == "cpu"
    np.testing.assert_allclose(out[0].double().numpy(), golden["n3_langevin/x_final_f32"], rtol=2e-5, atol=2e-6)
    np.testing.assert_allclose(out[0].double().numpy(), golden["n3_langevin/x_final_f64"], rtol=1e-4, atol=1e-5)
    imgs = fdsm.sampling.anneal_Langevin_dynamics(x0, net, sig.numpy(), n_steps_each=2, step_lr=2e-3)   # Philox noise
    assert len(imgs) == 6 and all(torch.isfinite(i).all() for i in imgs)


def test_ema_helper_matches_reference(fdsm, golden):
    lin = torch.nn.Linear(7, 5)
    nets.init_from_numpy(lin, 4400)
    lin = lin.to(DEV)
    h = fdsm.sampling.EMAHelper(mu=0.9)
    h.register(lin)
    for i in range(3):
        with torch.no_grad():
            for j, p in enumerate(lin.parameters()):
                p.add_((nets.np_randn(4410 + 10 * i + j, *p.shape).float() * 0.1).to(DEV))
        h.update(lin)
    got = torch.cat([h.shadow[n].reshape(-1) for n, _ in lin.named_parameters()]).double().cpu().numpy()
    np.testing.assert_allclose(got, golden["n4_ema/shadow_f32"], rtol=1e-6, atol=1e-7)
    big_s, big_p = torch.randn(1 << 20, device=DEV), torch.randn(1 << 20, device=DEV)
    want = (1. - 0.999) * big_p + 0.999 * big_s
    assert torch.equal(fdsm.ops.ema_update(big_s.clone(), big_p, 0.999), want)


def test_deepebm_samplers_match_reference(fdsm):
    """DeepEBM/ESM/sampling_mdsm.py (Annealed_Langevin_E, Langevin, Langevin_E) behind the reference's signatures: the update
    runs on fdsm_langevin_step; against outputs of the unmodified reference on replayed normal draws."""
    import os
    from oracle.make_golden_sampler import energy_net, inputs
    g = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "deepebm_sampler.npz"))
    net = energy_net(torch.float32).to(DEV)
    x0, zs, T = inputs(torch.float32)
    x0, zs = x0.to(DEV), [z.to(DEV) for z in zs]
    xs, E = fdsm.sampling.Annealed_Langevin_E(net, x0.clone(), 0.1, T.copy(), 3, noises=zs)
    assert len(xs) == 2 and xs[0].device.type == "cpu" and E.shape == (10, 6)
    # the network runs on cuBLAS here and on MKL in the reference: fp32 rounding of the energy gradient
    np.testing.assert_allclose(xs[0].double().numpy(), g["annealed/x_mid_f32"], rtol=1e-5, atol=1e-6)
    np.testing.assert_allclose(xs[1].double().numpy(), g["annealed/x_final_f32"], rtol=1e-5, atol=1e-6)
    np.testing.assert_allclose(xs[1].double().numpy(), g["annealed/x_final_f64"], rtol=1e-4, atol=1e-5)
    np.testing.assert_allclose(E.astype(np.float64), g["annealed/E_f32"], rtol=1e-4, atol=1e-5)
    xl = fdsm.sampling.Langevin(net, x0.clone(), 4, 0.1, noises=zs[:4])
    np.testing.assert_allclose(xl.double().cpu().numpy(), g["langevin/x_final_f32"], rtol=1e-5, atol=1e-6)
    xe, Em = fdsm.sampling.Langevin_E(net, x0.clone(), 4, 0.1, noises=zs[:4])
    assert torch.equal(xe, xl) and Em.shape == (4, 10)
    xn, _ = fdsm.sampling.Langevin_E(net, x0.clone(), 2, 0.1, noise=False)
    assert torch.isfinite(xn).all()
    # in-kernel Philox noise: finite, and different from the drift-only path
    xp = fdsm.sampling.Langevin(net, x0.clone(), 2, 0.1)
    assert torch.isfinite(xp).all() and not torch.equal(xp, xn)
