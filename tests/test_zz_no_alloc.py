"""raftk_corr_pyramid_static — the entry point for hosts that must not allocate.

Who needs it: a caller that reaches the library with inputs and outputs only and no way to
obtain scratch at that moment — a custom-call handler inside an already-compiled XLA program with
no scratch operand, or a process whose framework pool owns the GPU's memory so that a cudaMalloc
issued from the library would stall the device or fail.  The variant keeps its scratch in a fixed
16 MiB array inside the library's device image instead.

What is checked, straight through the C ABI (ctypes; the test itself may allocate its inputs and
outputs, the entry point may not): on two calls in a row, with different shapes so that the
second call runs over what the first left in the shared array, the result equals that of
raftk_corr_pyramid (caller-provided scratch) within the tolerance the pyramid tests use
(|d| <= 1e-3 * max(|ref|, rms(ref))); and a problem larger than the array is refused with
RAFTK_ERR_STATIC_LIMIT before anything is launched.  The small second shape is also compared with the
CPU oracle directly (float64 contraction, same tolerance), so the check does not rest on the other entry point alone.

The file name sorts last on purpose: with ``pytest -x`` every other test has already run if this
one fails.  It does not depend on them (passes alone in a fresh process too).
"""

import ctypes

import pytest
import torch

from jax_raft_b200 import _lib

pytestmark = pytest.mark.gpu

RAFTK_ERR_STATIC_LIMIT = -7


def _levels(q, h, w, dev, fill=None):
    out = []
    for l in range(4):
        t = torch.empty((q, h >> l, w >> l), dtype=torch.float32, device=dev)
        if fill is not None:
            t.fill_(fill)
        out.append(t)
    return out


def _rel_err(got, ref):
    ref = ref.double()
    rms = ref.pow(2).mean().sqrt()
    return float(((got.double() - ref).abs() / torch.maximum(ref.abs(), rms)).max())


def _run_both(lib, b, h, w, d, seed, dev, against_oracle=False):
    gen = torch.Generator(device=dev).manual_seed(seed)
    f1 = torch.randn((b, h, w, d), device=dev, generator=gen) * 1.45
    f2 = torch.randn((b, h, w, d), device=dev, generator=gen) * 1.45
    nq = h * w
    stream = ctypes.c_void_p(torch.cuda.current_stream(dev).cuda_stream)

    # existing entry point, scratch from the caller
    ref = _levels(b * nq, h, w, dev)
    nbytes = lib.raftk_corr_pyramid_scratch_bytes(b, nq, h, w, d)
    assert 0 < nbytes <= _lib.STATIC_SCRATCH_BYTES
    scratch = torch.empty((nbytes,), dtype=torch.uint8, device=dev)
    rc = lib.raftk_corr_pyramid(f1.data_ptr(), f2.data_ptr(), _lib.ptr_array(ref), 4, b, nq, h, w, d,
                                scratch.data_ptr(), nbytes, stream)
    _lib.check("raftk_corr_pyramid", rc)

    # new entry point: inputs and outputs only
    got = _levels(b * nq, h, w, dev, fill=float("nan"))
    rc = lib.raftk_corr_pyramid_static(f1.data_ptr(), f2.data_ptr(), _lib.ptr_array(got), 4, b, nq, h, w, d, stream)
    _lib.check("raftk_corr_pyramid_static", rc)
    torch.cuda.synchronize(dev)
    for l in range(4):
        assert torch.isfinite(got[l]).all(), f"level {l} not fully written"
        assert _rel_err(got[l], ref[l]) <= 1e-3, (l, _rel_err(got[l], ref[l]))
    if against_oracle:
        import numpy as np

        from oracle import corr_oracle as O

        want = O.build_pyramid(f1.cpu().numpy(), f2.cpu().numpy(), dtype=np.float64)
        for l in range(4):
            assert _rel_err(got[l], torch.from_numpy(want[l][..., 0]).to(dev)) <= 1e-3, l


def test_static_scratch_variant_matches_on_two_calls_in_a_row(cuda, lib):
    with torch.cuda.device(cuda):
        # call 1: the Sintel shape, which nearly fills the array (14.4 of 16 MiB)
        _run_both(lib, 1, 55, 128, 256, seed=0, dev=cuda)
        # call 2, right after: a smaller, ragged problem over the first call's leftovers
        _run_both(lib, 2, 19, 24, 128, seed=1, dev=cuda, against_oracle=True)


def test_static_scratch_variant_refuses_what_does_not_fit(cuda, lib):
    b, h, w, d = 2, 55, 128, 256  # 28.8 MB of split operands > 16 MiB
    assert lib.raftk_corr_pyramid_scratch_bytes(b, h * w, h, w, d) > _lib.STATIC_SCRATCH_BYTES
    with torch.cuda.device(cuda):
        f = torch.zeros((b, h, w, d), device=cuda)
        lv = [torch.zeros((1,), device=cuda) for _ in range(4)]  # never touched: the call must not launch
        rc = lib.raftk_corr_pyramid_static(f.data_ptr(), f.data_ptr(), _lib.ptr_array(lv), 4, b, h * w, h, w, d,
                                           ctypes.c_void_p(torch.cuda.current_stream(cuda).cuda_stream))
        torch.cuda.synchronize(cuda)
    assert rc == RAFTK_ERR_STATIC_LIMIT
    assert b"16 MiB" in lib.raftk_error_string(rc)
    assert all(float(t.sum()) == 0.0 for t in lv)
