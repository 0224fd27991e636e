"""CPU suite: the InstanceNorm++ (+ ELU) oracle (oracle/norm.py) against outputs and autograd gradients of the unmodified
reference modules (tests/golden/inorm_plus.npz, made by oracle/make_golden_norm.py), and the torch op sequence the product
uses for CPU tensors / double backward against the same vectors."""
import os

import numpy as np
import pytest
import torch

from oracle import norm as onorm

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "inorm_plus.npz")
CASES = ["plain_8x8", "plain_nobias_5x7", "plain_16x16", "cond_8x8", "cond_nobias_4x4"]


def load_case(gold, name):
    x, gout, table, labels = (gold[f"{name}/{k}"] for k in ("x", "gout", "table", "labels"))
    C = x.shape[1]
    cond, bias = name.startswith("cond"), "nobias" not in name
    rows = table[labels] if cond else table[:1]
    gamma, alpha = rows[:, :C], rows[:, C:2 * C]
    beta = rows[:, 2 * C:] if bias else None
    return x, gout, table, labels, gamma, alpha, beta, cond, bias


def param_grads(gold, key, dgamma, dalpha, dbeta, table, labels, cond, bias):
    """(ours, reference) pairs of parameter gradients in the reference's own layout."""
    C = dgamma.shape[1]
    if cond:
        want = gold[key + "dtable"]
        got = np.zeros_like(want)
        per_sample = np.concatenate([dgamma, dalpha] + ([dbeta] if bias else []), axis=1)
        np.add.at(got, labels, per_sample)
        return [(got, want)]
    pairs = [(dgamma.sum(0), gold[key + "dgamma"]), (dalpha.sum(0), gold[key + "dalpha"])]
    if bias:
        pairs.append((dbeta.sum(0), gold[key + "dbeta"]))
    return pairs


@pytest.mark.parametrize("name", CASES)
@pytest.mark.parametrize("act", [False, True])
def test_oracle_matches_reference_modules(name, act):
    gold = np.load(GOLD)
    x, gout, table, labels, gamma, alpha, beta, cond, bias = load_case(gold, name)
    key = f"{name}/{'elu' if act else 'id'}/f64/"
    out, cache = onorm.forward(x, gamma, alpha, beta, act)
    dx, dg, da, db = onorm.backward(gout, cache)
    np.testing.assert_allclose(out, gold[key + "out"], rtol=1e-10, atol=1e-11)
    np.testing.assert_allclose(dx, gold[key + "dx"], rtol=1e-9, atol=1e-10)
    for got, want in param_grads(gold, key, dg, da, db, table, labels, cond, bias):
        np.testing.assert_allclose(got, want, rtol=1e-9, atol=1e-10)
    # the reference's own fp32 run sits within fp32 rounding of it (the bound the GPU test holds our kernels to)
    key32 = key.replace("f64", "f32")
    assert np.abs(gold[key32 + "out"] - out).max() <= 2e-5 * max(1.0, np.abs(out).max())
    assert np.abs(gold[key32 + "dx"] - dx).max() <= 2e-5 * max(1.0, np.abs(dx).max())


@pytest.mark.parametrize("name", CASES)
def test_product_op_sequence_matches_reference_on_cpu(fdsm, name):
    gold = np.load(GOLD)
    x, gout, table, labels, gamma, alpha, beta, cond, bias = load_case(gold, name)
    t = lambda a: None if a is None else torch.tensor(np.ascontiguousarray(a), dtype=torch.float64)
    xt = t(x).requires_grad_()
    out = fdsm.ops.inorm_plus_reference(xt, t(gamma), t(alpha), t(beta), act=True)
    out.backward(t(gout))
    key = f"{name}/elu/f64/"
    np.testing.assert_allclose(out.detach().numpy(), gold[key + "out"], rtol=1e-10, atol=1e-11)
    np.testing.assert_allclose(xt.grad.numpy(), gold[key + "dx"], rtol=1e-9, atol=1e-10)
