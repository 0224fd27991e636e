"""CPU suite: the conv networks of fdsm_b200.nets (plain torch modules, generic mode) against forward outputs of the
UNMODIFIED reference modules (tests/golden/baseline_nets.npz, written by oracle/make_golden_baseline.py: same
name-keyed deterministic init loaded into both)."""
import os

import numpy as np
import torch

from oracle import nets as onets
from oracle.make_golden_baseline import SEED

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
G = np.load(os.path.join(ROOT, "tests", "golden", "baseline_nets.npz"))


def _rel(a, b):
    return float(np.abs(a - b).max() / np.abs(b).max())


def test_ncsnv2_forward_matches_reference(fdsm):
    m = onets.init_by_name(fdsm.nets.NCSNv2(3, 8, 12), SEED)
    x, y = onets.np_rand(SEED + 6, 3, 3, 32, 32).float(), torch.tensor([0, 5, 11])
    assert _rel(m(x, y).detach().numpy(), G["ncsnv2"]) < 1e-5


def test_cond_refinenet_forward_matches_reference(fdsm):
    m = onets.init_by_name(fdsm.nets.CondRefineNetDilated(3, 8, 12), SEED)
    x, y = onets.np_rand(SEED + 6, 3, 3, 32, 32).float(), torch.tensor([0, 5, 11])
    assert _rel(m(x, y).detach().numpy(), G["ncsn"]) < 1e-5


def test_res18_quadratic_forward_matches_reference(fdsm):
    m = onets.init_by_name(fdsm.nets.Res18Quadratic(1, 8, 32), SEED)
    assert _rel(m(onets.np_rand(SEED + 7, 3, 1, 32, 32).float()).detach().numpy(), G["res18"]) < 1e-5


def test_parameter_counts_match_the_survey(fdsm):
    """SURVEY.md section 8a (counts measured on the reference's modules): Res18_Quadratic n_chan 128 = 64.17 M,
    NCSNv2 29.7 M, CondRefineNetDilated 30.1 M."""
    n = lambda m: sum(p.numel() for p in m.parameters())
    assert n(fdsm.nets.Res18Quadratic(1, 128, 32)) == 64_171_779
    assert n(fdsm.nets.NCSNv2(3, 128, 232)) == 29_694_083
    assert n(fdsm.nets.CondRefineNetDilated(3, 128, 10)) == 30_117_635
