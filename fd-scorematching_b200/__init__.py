"""fd-scorematching_b200: B200-native (sm_100a) FD score-matching hot path.

Import as ``fdsm_b200`` (see the ``fdsm_b200.py`` shim at the repo root; the directory name
contains a hyphen).  Public surface = the reference's loss-function signatures (``losses``), the
kernel-level ops (``ops``) and the fused energy MLPs (``mlp``).
"""
from . import _lib, losses, mlp, nets, ops, sampling  # noqa: F401
from .losses import (FLAGS, MLE_efficient_score_matching_conjugate, anneal_ESM_scorenet_VR,  # noqa: F401
                     anneal_sliced_score_estimation_vr, deepebm_exact_score_matching, dsm_fd, exact_score_matching,
                     mdsm_baseline, mdsm_tracetrick,
                     dsm_tracetrick_FD, efficient_score_matching_conjugate, esm_scorenet_VR, fd_gradient_penalty, fd_loss, mdsm_fd,
                     mdsm_fd_nop, ncsn_anneal_ESM_scorenet, ncsn_anneal_ESM_scorenet_VR,
                     single_efficient_score_matching, single_sliced_score_matching, sliced_VR_score_matching,
                     ssm, ssm_fd, ssm_fd_nop, ssm_vr)
from .ops import manual_seed, set_row_offset  # noqa: F401

__all__ = ["losses", "ops", "mlp", "fd_loss", "manual_seed", "set_row_offset", "FLAGS"]
