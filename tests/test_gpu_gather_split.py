"""raftk_gather_rows_split on one GPU: the bands are ordinary local buffers here (the two-GPU run over peer
memory is tests/test_gpu_multi.py).  Bit-exact: the gathered map is a copy and the [hi | mid] split is the one
raftk_altcorr_split writes."""

import ctypes

import pytest
import torch

pytestmark = pytest.mark.gpu

ERR_NULL, ERR_SHAPE, ERR_UNSUPPORTED, ERR_ALIGN = -1, -2, -4, -6  # include/raftk.h


def _call(lib, bands, starts, full, halves, b, h, w, d, rank=0, flags=None, stream=None):
    from jax_raft_b200 import _lib

    world = len(bands)
    ptrs = (ctypes.c_void_p * world)(*[t.data_ptr() if t is not None else None for t in bands])
    fl = (ctypes.c_void_p * world)(*[t.data_ptr() if t is not None else None for t in flags]) if flags is not None else None
    st = (ctypes.c_int * (world + 1))(*starts)
    return lib.raftk_gather_rows_split(ptrs, fl, rank, st, world, full.data_ptr() if full is not None else None,
                                       halves.data_ptr() if halves is not None else None, b, h, w, d,
                                       stream if stream is not None else _lib.current_stream_ptr(torch.device("cuda", 0)))


@pytest.mark.parametrize("b,h,w,d,world", [(1, 16, 24, 256, 2), (2, 23, 32, 128, 3), (1, 136, 240, 256, 8), (3, 5, 7, 4, 8), (1, 9, 8, 64, 1)])
def test_gather_split_matches_copy_and_split(lib, b, h, w, d, world):
    from jax_raft_b200 import _lib, sharding

    dev = torch.device("cuda", 0)
    g = torch.Generator(device=dev).manual_seed(h * w + world)
    f2 = torch.randn((b, h, w, d), device=dev, generator=g) * 1.7
    starts = [sharding.row_range(h, world, k)[0] for k in range(world)] + [h]
    bands = [f2[:, starts[k]:starts[k + 1]].contiguous() for k in range(world)]  # h < world: some bands are empty
    full = torch.full((b, h, w, d), float("nan"), device=dev)
    halves = torch.zeros((b, h, w, 2 * d), dtype=torch.bfloat16, device=dev)
    assert _call(lib, bands, starts, full, halves, b, h, w, d) == 0
    ref = torch.empty_like(halves)
    assert lib.raftk_altcorr_split(f2.data_ptr(), ref.data_ptr(), b * h * w, d, 1.0, _lib.current_stream_ptr(dev)) == 0
    torch.cuda.synchronize()
    assert torch.equal(full, f2)
    assert torch.equal(halves.view(torch.int16), ref.view(torch.int16))
    # gather only; the walk over the bands may start anywhere
    for rank in range(world):
        full.fill_(float("nan"))
        assert _call(lib, bands, starts, full, None, b, h, w, d, rank=rank) == 0
        torch.cuda.synchronize()
        assert torch.equal(full, f2)


def test_gather_split_unbalanced_bands(lib):
    """Any non-decreasing row_start is accepted: empty bands in the middle, the largest band last."""
    dev = torch.device("cuda", 0)
    b, h, w, d = 2, 11, 9, 32
    f2 = torch.randn((b, h, w, d), device=dev)
    for starts in ([0, 2, 2, 3, 11], [0, 0, 11], [0, 11, 11, 11], [0, 1, 5, 6, 7, 8, 9, 10, 11]):
        bands = [f2[:, starts[k]:starts[k + 1]].contiguous() for k in range(len(starts) - 1)]
        full = torch.full_like(f2, float("nan"))
        halves = torch.zeros((b, h, w, 2 * d), dtype=torch.bfloat16, device=dev)
        assert _call(lib, bands, starts, full, halves, b, h, w, d, rank=len(bands) - 1) == 0
        torch.cuda.synchronize()
        assert torch.equal(full, f2)
        hi = halves[..., :d].float()
        assert torch.equal(hi, f2.bfloat16().float())
        assert torch.equal(halves[..., d:].float(), (f2 - hi).bfloat16().float())


def test_gather_split_in_kernel_sync_two_ranks_one_gpu(lib):
    """The flag protocol with both "ranks" on this GPU, each on its own stream (grids small enough to be
    co-resident): five calls each, results right every time, epochs in step, no wait gave up."""
    dev = torch.device("cuda", 0)
    b, h, w, d = 1, 8, 8, 64
    starts = [0, 4, 8]
    bands = [torch.empty((b, 4, w, d), device=dev) for _ in range(2)]
    flags = [torch.zeros((64,), dtype=torch.int32, device=dev) for _ in range(2)]
    streams = [torch.cuda.Stream(dev) for _ in range(2)]
    fulls = [torch.empty((b, h, w, d), device=dev) for _ in range(2)]
    torch.cuda.synchronize()
    for rep in range(5):
        f2 = torch.randn((b, h, w, d), device=dev)
        torch.cuda.synchronize()
        for r in (rep % 2, 1 - rep % 2):  # either rank may be launched first
            with torch.cuda.stream(streams[r]):
                bands[r].copy_(f2[:, starts[r]:starts[r + 1]])
                assert _call(lib, bands, starts, fulls[r], None, b, h, w, d, rank=r, flags=flags, stream=streams[r].cuda_stream) == 0
        torch.cuda.synchronize()
        assert torch.equal(fulls[0], f2) and torch.equal(fulls[1], f2)
    for r in range(2):
        fl = flags[r].tolist()
        assert fl[16] == 5 and fl[17] == 0 and fl[18] == 0, fl[:19]  # epoch, blocks_done, timeout
        assert fl[1 - r] == 5 and fl[8 + 1 - r] == 5                 # arrive / done written by the other rank


def test_gather_split_rejects(lib):
    dev = torch.device("cuda", 0)
    b, h, w, d = 1, 8, 8, 64
    f2 = torch.zeros((b, h, w, d), device=dev)
    full = torch.empty_like(f2)
    bands = [f2[:, :4].contiguous(), f2[:, 4:].contiguous()]
    assert _call(lib, bands, [0, 4, 8], None, None, b, h, w, d) == ERR_NULL
    assert _call(lib, [bands[0], None], [0, 4, 8], full, None, b, h, w, d) == ERR_NULL
    assert _call(lib, bands, [0, 4, 7], full, None, b, h, w, d) == ERR_SHAPE   # bands do not cover h
    assert _call(lib, bands, [0, 4, 8], full, None, b, h, w, 0) == ERR_SHAPE
    assert _call(lib, bands, [0, 6, 4], full, None, b, h, w, d) == ERR_SHAPE  # decreasing / wrong end
    assert _call(lib, bands * 5, [0] * 10 + [8], full, None, b, h, w, d) == ERR_SHAPE  # more than 8 ranks
    assert _call(lib, bands, [0, 4, 8], full, None, b, h, w, 6) == ERR_UNSUPPORTED  # D % 4
    odd = torch.empty((b * h * w * d + 4,), device=dev)[1:1 + b * h * w * d]
    assert _call(lib, bands, [0, 4, 8], odd, None, b, h, w, d) == ERR_ALIGN
    assert _call(lib, bands, [0, 4, 8], full, None, 0, h, w, d) == 0  # empty batch: nothing to do
    flags = [torch.zeros((64,), dtype=torch.int32, device=dev) for _ in range(2)]
    assert _call(lib, bands, [0, 4, 8], full, None, b, h, w, d, rank=2, flags=flags) == ERR_SHAPE
    assert _call(lib, bands, [0, 4, 8], full, None, b, h, w, d, rank=0, flags=[flags[0], None]) == ERR_NULL
    assert _call(lib, bands, [0, 4, 8], full, None, b, h, w, d, rank=0, flags=[flags[0], flags[1][1:]]) == ERR_ALIGN
