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
