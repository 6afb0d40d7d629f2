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


def _stream(cuda):
    return ctypes.c_void_p(torch.cuda.current_stream(cuda).cuda_stream)


@pytest.mark.parametrize("hid,extra", [(128, 256), (96, 146), (6, 10)])
def test_gru_gate_and_update(cuda, lib, hid, extra):
    """ConvGRU's element-wise halves (model.py:306-311) on channel slices of the conv input buffers."""
    gen = torch.Generator(device=cuda).manual_seed(hid)
    npix, ld = 2 * 7 * 9, hid + extra
    A = torch.randn((npix, ld), device=cuda, generator=gen)
    B = torch.randn((npix, ld), device=cuda, generator=gen)
    zr_pre = torch.randn((npix, 2 * hid), device=cuda, generator=gen) * 2
    bzr = torch.randn((2 * hid,), device=cuda, generator=gen)
    q_pre = torch.randn((npix, hid), device=cuda, generator=gen) * 2
    bq = torch.randn((hid,), device=cuda, generator=gen)
    h = A[:, :hid].clone()
    z_ref = torch.sigmoid(zr_pre[:, :hid] + bzr[:hid])
    rh_ref = torch.sigmoid(zr_pre[:, hid:] + bzr[hid:]) * h
    h_ref = (1 - z_ref) * h + z_ref * torch.tanh(q_pre + bq)
    A0, B0 = A.clone(), B.clone()

    z = torch.empty((npix, hid), device=cuda)
    with torch.cuda.device(cuda):
        rc = lib.raftk_gru_gate(zr_pre.data_ptr(), bzr.data_ptr(), A.data_ptr(), ld, z.data_ptr(), B.data_ptr(), ld, npix, hid,
                                _stream(cuda))
    _lib.check("raftk_gru_gate", rc)
    assert torch.allclose(z, z_ref, rtol=1e-6, atol=1e-6)
    assert torch.allclose(B[:, :hid], rh_ref, rtol=1e-6, atol=1e-6)
    assert torch.equal(B[:, hid:], B0[:, hid:]) and torch.equal(A, A0)  # nothing outside the slice moves

    hd = torch.empty((npix, hid), device=cuda)
    with torch.cuda.device(cuda):
        rc = lib.raftk_gru_update(q_pre.data_ptr(), bq.data_ptr(), z.data_ptr(), A.data_ptr(), ld, hd.data_ptr(), npix, hid,
                                  _stream(cuda))
    _lib.check("raftk_gru_update", rc)
    assert torch.allclose(hd, h_ref, rtol=1e-5, atol=1e-6)
    assert torch.equal(A[:, :hid], hd) and torch.equal(A[:, hid:], A0[:, hid:])


@pytest.mark.parametrize("c_cf,ld_src,off,ld", [(126, 126, 256, 384), (126, 128, 256, 384), (80, 80, 160, 242)])
def test_motion_scatter(cuda, lib, c_cf, ld_src, off, ld):
    """MotionEncoder tail (model.py:286-290): relu(conv + bias) ++ flow into both GRU input buffers."""
    gen = torch.Generator(device=cuda).manual_seed(c_cf)
    npix = 3 * 5 * 11
    cf_full = torch.randn((npix, ld_src), device=cuda, generator=gen)  # channels >= c_cf: zero-filter padding
    cf_pre = cf_full[:, :c_cf]
    bias = torch.randn((ld_src,), device=cuda, generator=gen)[:c_cf]
    flow = torch.randn((npix, 2), device=cuda, generator=gen)
    A = torch.full((npix, ld), 7.0, device=cuda)
    B = torch.full((npix, ld), 9.0, device=cuda)
    with torch.cuda.device(cuda):
        rc = lib.raftk_motion_scatter(cf_full.data_ptr(), ld_src, bias.data_ptr(), flow.data_ptr(), A.data_ptr() + 4 * off,
                                      B.data_ptr() + 4 * off, ld, npix, c_cf, _stream(cuda))
    _lib.check("raftk_motion_scatter", rc)
    ref = torch.cat([F.relu(cf_pre + bias), flow], dim=1)
    for buf, fill in ((A, 7.0), (B, 9.0)):
        assert torch.equal(buf[:, off:off + c_cf + 2], ref)
        assert (buf[:, :off] == fill).all() and (buf[:, off + c_cf + 2:] == fill).all()


def test_gru_glue_rejects_bad_shapes(cuda, lib):
    x = torch.zeros((64,), device=cuda)
    p = x.data_ptr()
    assert lib.raftk_gru_gate(p, p, p, 7, p, p, 8, 1, 4, None) == -2      # odd stride
    assert lib.raftk_gru_gate(p, p, p, 2, p, p, 8, 1, 4, None) == -2      # stride < hidden
    assert lib.raftk_gru_update(p, p, p, p + 4, 8, p, 1, 4, None) == -6   # 4-byte aligned pointer
    assert lib.raftk_motion_scatter(p, 8, p, p, p, p, 8, 1, 8, None) == -2   # ld < c_cf + 2
    assert lib.raftk_motion_scatter(p, 6, p, p, p, p, 16, 1, 8, None) == -2  # source pitch < c_cf
    assert lib.raftk_flow_update(p, 3, p, p, p, p, None, 0, 1, None) == -2   # odd delta pitch
    assert lib.raftk_gru_gate(None, p, p, 8, p, p, 8, 1, 4, None) == -1
    assert lib.raftk_gru_gate(None, None, None, 8, None, None, 8, 0, 4, None) == 0  # empty batch


@pytest.mark.parametrize("ld_delta,ld_flow", [(2, 0), (8, 8)])
def test_flow_update(cuda, lib, ld_delta, ld_flow):
    """UpdateCell tail (model.py:504-506), bit-exact against the same float32 sequence in torch; with a
    padded flow-head output and the second, pitched copy of the flow."""
    gen = torch.Generator(device=cuda).manual_seed(5)
    npix = 2 * 13 * 17
    delta_full = torch.randn((npix, ld_delta), device=cuda, generator=gen)
    delta = delta_full[:, :2]
    bias = torch.randn((8,), device=cuda, generator=gen)[:2]
    c0 = torch.rand((npix, 2), device=cuda, generator=gen) * 100
    c1 = c0 + torch.randn((npix, 2), device=cuda, generator=gen) * 5
    ref_c1 = c1 + (delta + bias)
    ref_flow = ref_c1 - c0
    flow = torch.empty_like(c1)
    pitched = torch.full((npix, max(ld_flow, 1)), 3.0, device=cuda)
    with torch.cuda.device(cuda):
        rc = lib.raftk_flow_update(delta_full.data_ptr(), ld_delta, bias.data_ptr(), c0.data_ptr(), c1.data_ptr(), flow.data_ptr(),
                                   pitched.data_ptr() if ld_flow else None, ld_flow, npix, _stream(cuda))
    _lib.check("raftk_flow_update", rc)
    assert torch.equal(c1, ref_c1) and torch.equal(flow, ref_flow)
    if ld_flow:
        assert torch.equal(pitched[:, :2], ref_flow) and (pitched[:, 2:] == 3.0).all()


@pytest.mark.parametrize("shape", [(2, 55, 128, 128), (1, 23, 31, 8)])
@pytest.mark.parametrize("relu", [0, 1])
def test_instance_norm_with_residual_tail(cuda, lib, shape, relu):
    """raftk_instance_norm_residual_nhwc: relu(residual + [relu](instance_norm(x))) (model.py:183-184)."""
    n, h, w, c = shape
    gen = torch.Generator(device=cuda).manual_seed(c + relu)
    x = torch.randn(shape, device=cuda, generator=gen) * 2.5 + 10.0
    res = torch.randn(shape, device=cuda, generator=gen)
    y = F.instance_norm(x.permute(0, 3, 1, 2).double(), eps=1e-5).permute(0, 2, 3, 1)
    if relu:
        y = F.relu(y)
    ref = F.relu(res.double() + y)
    got = x.clone()
    nbytes = lib.raftk_instance_norm_scratch_bytes(n, c)
    scratch = torch.empty((nbytes,), dtype=torch.uint8, device=cuda)
    with torch.cuda.device(cuda):
        rc = lib.raftk_instance_norm_residual_nhwc(got.data_ptr(), res.data_ptr(), n, h * w, c, 1e-5, relu, scratch.data_ptr(),
                                                   nbytes, _stream(cuda))
    _lib.check("raftk_instance_norm_residual_nhwc", rc)
    assert torch.allclose(got.double(), ref, rtol=1e-4, atol=1e-4)


def test_add_relu(cuda, lib):
    gen = torch.Generator(device=cuda).manual_seed(1)
    x = torch.randn((3, 17, 19, 8), device=cuda, generator=gen)
    y = torch.randn((3, 17, 19, 8), device=cuda, generator=gen)
    ref = F.relu(x + y)
    with torch.cuda.device(cuda):
        rc = lib.raftk_add_relu(x.data_ptr(), y.data_ptr(), x.numel(), _stream(cuda))
    _lib.check("raftk_add_relu", rc)
    assert torch.equal(x, ref)
    assert lib.raftk_add_relu(x.data_ptr(), y.data_ptr(), 6, None) == -2
