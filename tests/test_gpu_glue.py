"""The two host-model glue kernels (not hot-path rows): bias + activation and InstanceNorm(+ReLU)
on NHWC conv outputs, against the torch ops they replace."""

import ctypes

import pytest
import torch
import torch.nn.functional as F

from jax_raft_b200 import _lib

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("c", [2, 64, 126, 256])
@pytest.mark.parametrize("act", [0, 1, 2, 3])
def test_bias_act(cuda, lib, c, act):
    gen = torch.Generator(device=cuda).manual_seed(c + act)
    x = torch.randn((3, 11, 13, c), device=cuda, generator=gen) * 3
    b = torch.randn((c,), device=cuda, generator=gen)
    ref = x + b
    ref = [ref, F.relu(ref), torch.sigmoid(ref), torch.tanh(ref)][act]
    y = x.clone()
    with torch.cuda.device(cuda):
        rc = lib.raftk_bias_act_nhwc(y.data_ptr(), b.data_ptr(), 3 * 11 * 13, c, act,
                                     ctypes.c_void_p(torch.cuda.current_stream(cuda).cuda_stream))
    _lib.check("raftk_bias_act_nhwc", rc)
    assert torch.allclose(y, ref, rtol=1e-6, atol=1e-6)


@pytest.mark.parametrize("shape", [(2, 55, 128, 128), (2, 110, 256, 96), (1, 23, 31, 8), (4, 220, 512, 64)])
@pytest.mark.parametrize("relu", [0, 1])
def test_instance_norm(cuda, lib, shape, relu):
    n, h, w, c = shape
    gen = torch.Generator(device=cuda).manual_seed(c)
    x = torch.randn(shape, device=cuda, generator=gen) * 2.5 + 40.0  # large mean: one-pass variance must not cancel
    ref = F.instance_norm(x.permute(0, 3, 1, 2).double(), eps=1e-5).permute(0, 2, 3, 1)
    if relu:
        ref = F.relu(ref)
    y = x.clone()
    nbytes = lib.raftk_instance_norm_scratch_bytes(n, c)
    scratch = torch.empty((nbytes,), dtype=torch.uint8, device=cuda)
    with torch.cuda.device(cuda):
        rc = lib.raftk_instance_norm_nhwc(y.data_ptr(), n, h * w, c, 1e-5, relu, scratch.data_ptr(), nbytes,
                                          ctypes.c_void_p(torch.cuda.current_stream(cuda).cuda_stream))
    _lib.check("raftk_instance_norm_nhwc", rc)
    assert torch.allclose(y.double(), ref, rtol=0, atol=2e-5)
