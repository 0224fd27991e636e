"""fd-scorematching_b200: B200-native (sm_100a) FD score-matching hot path.

Import as ``fdsm_b200`` (see the ``fdsm_b200.py`` shim at the repo root; the directory name
contains a hyphen).  Public surface = the reference's loss-function signatures (``losses``), the
kernel-level ops (``ops``) and the fused energy MLPs (``mlp``).
"""
import os as _os

# The end-to-end step keeps several streams busy at once (two H2D copy streams, the step, the weight-gradient side streams).
# With the CUDA default of 8 hardware queues, two of them can land on one queue and serialise (measured on B200: the fp32 e2e
# step then takes copy + step = 380 us instead of the copy's 236 us, in some processes and not in others).  Read by the driver
# when the context is created, so it only has an effect when this package is imported before CUDA is initialised.
_os.environ.setdefault("CUDA_DEVICE_MAX_CONNECTIONS", "32")

from . import _lib, losses, mlp, nets, ops, sampling  # noqa: F401,E402
from .losses import (FLAGS, MLE_efficient_score_matching_conjugate, anneal_ESM_scorenet_VR,  # noqa: F401
                     anneal_sliced_score_estimation_vr, deepebm_exact_score_matching, dsm_fd, exact_score_matching,
                     mdsm_baseline, mdsm_tracetrick,
                     dsm_tracetrick_FD, efficient_score_matching_conjugate, esm_scorenet_VR, fd_gradient_penalty, fd_loss, mdsm_fd,
                     mdsm_fd_nop, ncsn_anneal_ESM_scorenet, ncsn_anneal_ESM_scorenet_VR,
                     single_efficient_score_matching, single_sliced_score_matching, sliced_VR_score_matching,
                     ssm, ssm_fd, ssm_fd_nop, ssm_vr)
from .ops import manual_seed, set_row_offset  # noqa: F401

__all__ = ["losses", "ops", "mlp", "fd_loss", "manual_seed", "set_row_offset", "FLAGS"]
