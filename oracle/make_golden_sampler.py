"""Golden outputs of the UNMODIFIED reference's DeepEBM sampler (DeepEBM/ESM/sampling_mdsm.py: Annealed_Langevin_E, Langevin) on a
small quadratic-head MLP energy, with the normal draws replayed from numpy streams -> tests/golden/deepebm_sampler.npz.
Runs in the build container only (/root/reference is not on the GPU box).  TEST INFRASTRUCTURE.
    python -m oracle.make_golden_sampler"""
import importlib.util
import os

import numpy as np
import torch

from . import nets, sampling as osamp
from .make_golden import REF, Replay

OUT = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden")
SEED = 5150


def energy_net(dt):
    net = nets.QuadraticMLP(24, 16, 2, act="elu")
    nets.init_from_numpy(net, SEED)
    return net.to(dt)


def inputs(dt):
    x0 = nets.np_rand(SEED + 1, 10, 24).to(dt)
    zs = [nets.np_randn(SEED + 10 + i, 10, 24).to(dt) for i in range(6)]
    T = np.linspace(1.0, 0.2, 6)
    return x0, zs, T


def main():
    spec = importlib.util.spec_from_file_location("ref_sampling_mdsm", f"{REF}/DeepEBM/ESM/sampling_mdsm.py")
    ref = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(ref)
    out = {}
    for tag, dt in (("f32", torch.float32), ("f64", torch.float64)):
        net = energy_net(dt)
        x0, zs, T = inputs(dt)
        with Replay([z.clone() for z in zs]):
            xs, E = ref.Annealed_Langevin_E(net, x0.clone(), 0.1, T.copy(), 3)
        out[f"annealed/x_mid_{tag}"], out[f"annealed/x_final_{tag}"] = xs[0].double().numpy(), xs[1].double().numpy()
        out[f"annealed/E_{tag}"] = np.asarray(E, dtype=np.float64)
        mine, Em = osamp.mdsm_annealed_langevin(net, x0.clone(), 0.1, T, zs)
        print(f"annealed {tag}: |ref - oracle| max x {float((xs[1] - mine).abs().max()):.2e}  E {float(np.abs(E - Em.numpy()).max()):.2e}")
        with Replay([z.clone() for z in zs[:4]]):
            xl = ref.Langevin(net, x0.clone(), 4, 0.1)
        out[f"langevin/x_final_{tag}"] = xl.detach().double().numpy()
        mine, _ = osamp.mdsm_annealed_langevin(net, x0.clone(), 0.1, np.ones(4), zs[:4])
        print(f"langevin {tag}: |ref - oracle| max x {float((xl.detach() - mine).abs().max()):.2e}")
    np.savez_compressed(os.path.join(OUT, "deepebm_sampler.npz"), **out)
    print("wrote", os.path.join(OUT, "deepebm_sampler.npz"))


if __name__ == "__main__":
    main()
