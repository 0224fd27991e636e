"""Stage-by-stage check of the chain kernel (fdsm_jet_fd_ssm) against the float64 jet oracle: inputs XV, every layer's
(h0, a, b), every dZ_l, loss, per-tensor gradients.  Run on the GPU box:
    python tests/tools/jet_debug.py [D H L act head B passes]"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import fdsm_b200 as fdsm  # noqa: E402
from oracle import fd_losses, jet, nets  # noqa: E402


def rel(a, b):
    a, b = a.detach().double().cpu(), b.detach().double().cpu()
    return float((a - b).norm() / b.norm().clamp_min(1e-300))


def main(D=64, H=48, L=3, act="nsoftplus", head="quadratic", B=37, passes=2, eps=0.05):
    dev = "cuda:0"
    ref = (nets.QuadraticMLP(D, H, L, act) if head == "quadratic" else nets.SoftplusMLP(D, H))
    nets.init_from_numpy(ref, 11)
    ref = ref.double()
    net = fdsm.mlp.FusedEnergyMLP(D, H, L, act=act, head=head, device=dev, formulation="jet", jet_passes=passes)
    net.load_reference_state_dict([p.detach().float() for p in ref.parameters()])
    x, u = nets.np_rand(1, B, D), nets.np_randn(2, B, D)
    v = fd_losses.sphere_direction(u, eps)
    scale = 1.0 / eps ** 2
    loss, grad = net.fd_ssm_loss_and_grad(x.float().to(dev), eps, scale, u=u.float().to(dev))
    torch.cuda.synchronize()
    st = net._step_state
    o = st["offsets"]
    S, Bp, npad, kpad0, kpadh, xv = o[0], o[1], o[2], o[3], o[4], o[5]
    ws = st["ws"]
    print(f"S={S} Bp={Bp} npad={npad} kpad0={kpad0} kpadh={kpadh} tiles={o[7]} passes={o[6]} splits={[o[11 + 4 * l] for l in range(L)]}")
    XV = ws[xv:xv + 2 * Bp * kpad0].view(2, Bp, kpad0)
    print("XV x  ", rel(XV[0, :B, :D], x), " ones col", float(XV[0, :B, D].min()), float(XV[0, :B, D].max()), " pad max", float(XV[0, :, D + 1:].abs().max()))
    print("XV v  ", rel(XV[1, :B, :D], v), " pad max", float(XV[1, :, D:].abs().max()))
    if hasattr(ref, "ln1"):
        layers = [(getattr(ref, f"fc{i + 1}").weight, getattr(ref, f"fc{i + 1}").bias) for i in range(L)]
    else:
        lins = [m for m in ref.net if isinstance(m, torch.nn.Linear)]
        layers = [(m.weight, m.bias) for m in lins[:-1]]
    pre, post = jet.mlp_jet_trace(layers, act if hasattr(ref, "ln1") else "softplus", x, v)
    if hasattr(ref, "ln1"):
        E0, dE, sE = jet.quadratic_head_jet(ref.ln1.weight.view(-1), ref.ln1.bias, ref.ln2.weight.view(-1), ref.ln2.bias,
                                            ref.lq.weight.view(-1), ref.lq.bias, *post[-1])
    else:
        w, bias = lins[-1].weight.view(-1), lins[-1].bias
        E0, dE, sE = post[-1][0] @ w + bias, 2.0 * (post[-1][1] @ w), 2.0 * (post[-1][2] @ w)
    want = (dE * dE / 8.0 - sE).mean() * scale
    want.backward()
    print("loss  ", float(loss), float(want), abs(float(loss) - float(want)) / abs(float(want)))
    for l in range(1, L):
        h = ws[o[9 + 4 * l]:o[9 + 4 * l] + 3 * Bp * kpadh].view(3, Bp, kpadh)
        print(f"H_{l}   h0 {rel(h[0, :B, :H], post[l - 1][0]):.2e}  a {rel(h[1, :B, :H], post[l - 1][1]):.2e}  b {rel(h[2, :B, :H], post[l - 1][2]):.2e}"
              f"  ones {float(h[0, :B, H].min())} pad {float(h[:, :, H + 1:].abs().max())} {float(h[1:, :, H].abs().max())}")
    for l in range(L - 1, -1, -1):
        ns = 3 if l else 2
        dz = ws[o[10 + 4 * l]:o[10 + 4 * l] + ns * Bp * npad].view(ns, Bp, npad)
        names = ["dz0", "dP", "dQ"]
        print(f"dZ_{l}  " + "  ".join(f"{names[s]} {rel(dz[s, :B, :H], pre[l][s].grad):.2e}" for s in range(ns)) +
              f"  pad {float(dz[:, :, H:].abs().max()) if npad > H else 0.0} tail {float(dz[:, B:].abs().max()) if Bp > B else 0.0}")
    g64 = torch.cat([(p.grad if p.grad is not None else torch.zeros_like(p)).reshape(-1) for p in ref.parameters()])
    print("grad  ", rel(grad, g64))
    views = net.named_views(grad)
    for (name, _), p in zip(net.param_shapes(), ref.parameters()):
        if p.grad is not None:
            print(f"   {name:12s} {rel(views[name].reshape(-1), p.grad.reshape(-1)):.2e}")
        else:
            print(f"   {name:12s} |ours| = {float(views[name].abs().max()):.2e} (no gradient in the oracle: c1 = 2 drops E0)")


if __name__ == "__main__":
    a = sys.argv[1:]
    if a:
        main(int(a[0]), int(a[1]), int(a[2]), a[3], a[4], int(a[5]), int(a[6]))
    else:
        main()
