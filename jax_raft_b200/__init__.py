"""jax_raft_b200 — B200 (sm_100a) implementation of RAFT's correlation hot path behind the
API of alebeck/jax-raft (``jax_raft/__init__.py:1-7``: ``RAFT``, ``raft_large``, ``raft_small``).

Importing the package does not need a GPU; calling any correlation / upsampling op does, and
raises if ``lib/libraftk.so`` has not been built (there is no CPU or eager fallback).
"""

from .corr import AltCorrBlock, CorrBlock
from .functional import (grid_sample, make_coords_grid, mask_upsample, pack_mask_predictor, resize_with_aligned_corners,
                         upsample_flow)
from .model import (RAFT, RAFT_LARGE, RAFT_SMALL, RaftConfig, init_variables, raft_large, raft_small,
                    torch_state_dict_from_variables, variables_from_torch_state_dict)

__all__ = [
    "RAFT", "raft_large", "raft_small", "CorrBlock", "AltCorrBlock", "upsample_flow", "mask_upsample", "pack_mask_predictor", "make_coords_grid", "grid_sample", "resize_with_aligned_corners",
    "RaftConfig", "RAFT_LARGE", "RAFT_SMALL", "init_variables", "variables_from_torch_state_dict",
    "torch_state_dict_from_variables",
]
