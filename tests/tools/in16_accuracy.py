"""Loss / gradient error of the chain kernel against the float64 oracle at config 2 (B = 4096, D = 784, 200 x 3, nsoftplus) for the
weight formats jet_passes = 3 / 2 / 4 (4 = input layer in kind::f16).  TEST TOOL (imports the oracle).
    python tests/tools/in16_accuracy.py"""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import fdsm_b200 as fdsm  # noqa: E402
from oracle import fd_losses, nets  # noqa: E402
import test_jet_gpu as T  # noqa: E402

D, H, L, B, eps = 784, 200, 3, 4096, 0.05
for data, mk in (("uniform [0,1) pixels", lambda: nets.np_rand(1, B, D)),
                 ("uniform pixels + 0.005 N(0,1) noise (train_ssm.py:62)", lambda: nets.np_rand(1, B, D) + 0.005 * nets.np_randn(5, B, D))):
    x, u = mk(), nets.np_randn(2, B, D)
    print(data)
    for passes in (3, 2, 4):
        net, ref = T._pair(fdsm, D, H, L, "nsoftplus", "quadratic", passes=passes)
        want, gwant = T._oracle(ref, x, fd_losses.sphere_direction(u, eps), 1.0 / eps ** 2)
        loss, grad = net.fd_ssm_loss_and_grad(x.float().to(T.DEV), eps, 1.0 / eps ** 2, u=u.float().to(T.DEV))
        print(f"   jet_passes = {passes}: loss rel err {abs(float(loss) - want) / abs(want):.2e}   gradient rel err (L2) {T._relv(grad, gwant):.2e}")
