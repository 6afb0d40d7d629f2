"""Host-side mirror of the reference API: variable tree, converter rules, argument checks.  No GPU."""

import numpy as np
import pytest
import torch

import jax_raft_b200 as J
from jax_raft_b200.model import _tree_items


@pytest.mark.parametrize("name", ["raft_large", "raft_small"])
def test_variable_tree_matches_converted_torchvision_checkpoint(name):
    import torchvision.models.optical_flow as tvof

    model, variables = getattr(J, name)()
    assert set(variables) == {"params", "batch_stats"}
    tv = getattr(tvof, name)(weights=None)
    conv = J.variables_from_torch_state_dict(tv.state_dict())  # convert_checkpoint.py rules
    mine = dict(_tree_items(variables))
    ref = dict(_tree_items(conv))
    assert set(mine) == set(ref)
    for k in mine:
        assert np.asarray(mine[k]).shape == np.asarray(ref[k]).shape, k
    # e.g. SURVEY.md §8c: (256,324,1,1) -> params/update_block/motion_encoder/convcorr1/layers_0/kernel (1,1,324,256)
    if name == "raft_large":
        k = variables["params"]["update_block"]["motion_encoder"]["convcorr1"]["layers_0"]["kernel"]
        assert k.shape == (1, 1, 324, 256)
        assert "mean" in variables["batch_stats"]["context_encoder"]["convnormrelu"]["layers_1"]
    # round trip
    sd = {k: v for k, v in tv.state_dict().items() if not k.endswith("num_batches_tracked")}
    back = J.torch_state_dict_from_variables(conv)
    assert set(back) == set(sd)
    assert all(torch.equal(back[k], sd[k]) for k in sd)


def test_factories_keep_reference_signature():
    model, variables = J.raft_large(pretrained=False)
    assert model.corr_block.out_channels == 4 * 81 and model.corr_block.radius == 4
    small, _ = J.raft_small()
    assert small.corr_block.out_channels == 4 * 49 and small.config.use_mask_predictor is False
    with pytest.raises(RuntimeError, match="network"):
        J.raft_large(pretrained=True)
    with pytest.raises(TypeError):
        J.raft_large(not_a_kwarg=1)
    alt = J.AltCorrBlock(4, 4)
    m2, _ = J.raft_large(corr_block=alt)  # the reference's injection point, model.py:643-644
    assert m2.corr_block is alt
    with pytest.raises(AssertionError):
        J.raft_large(corr_block=J.CorrBlock(4, 3))  # out_channels must match the motion encoder
    with pytest.raises(NotImplementedError):
        model.apply(variables, np.zeros((1, 64, 64, 3), np.float32), np.zeros((1, 64, 64, 3), np.float32), train=True)
    # host-side options (not in the reference): scheduling / fusion switches, all consistent with each other
    assert model.fused_update and model.concurrent and not model.fused_lookup_conv
    plain, _ = J.raft_large(fused_update=False, fused_lookup_conv=True)
    assert not plain.concurrent and not plain.fused_lookup_conv  # both need the fused update cell
    fp32, _ = J.raft_large(precision="fp32", fused_lookup_conv=True)
    assert not fp32.fused_lookup_conv                           # the fused kernel multiplies in TF32
    model._prepared["x"] = model._graphs["y"] = object()
    model.invalidate()
    assert not model._prepared and not model._graphs


def test_corr_block_asserts_like_reference_and_has_no_cpu_fallback():
    cb = J.CorrBlock(4, 4)
    assert cb.out_channels == 324
    with pytest.raises(AssertionError, match="same shapes"):
        cb.build_pyramid(torch.zeros(1, 16, 16, 8), torch.zeros(1, 16, 17, 8))
    with pytest.raises(AssertionError, match="too small"):
        cb.build_pyramid(torch.zeros(1, 15, 16, 8), torch.zeros(1, 15, 16, 8))
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        cb.build_pyramid(torch.zeros(1, 16, 16, 8), torch.zeros(1, 16, 16, 8))
    with pytest.raises(AssertionError):
        cb.index_pyramid([torch.zeros(4, 16, 16, 1)] * 4, torch.zeros(1, 2, 2, 3))
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        J.upsample_flow(torch.zeros(1, 4, 4, 2), torch.zeros(1, 4, 4, 576))
    with pytest.raises(AssertionError):
        J.upsample_flow(torch.zeros(1, 4, 4, 2), torch.zeros(1, 4, 4, 575))


def test_bilinear_branch_and_coords_grid_on_cpu():
    from oracle import corr_oracle as O

    flow = torch.randn(2, 5, 7, 2)
    up = J.upsample_flow(flow)  # raft_small path: host-framework code (model.py:79-83)
    np.testing.assert_allclose(up.numpy(), O.upsample_flow(flow.numpy()), rtol=1e-4, atol=1e-4)
    g = J.make_coords_grid(2, 3, 4, device="cpu")
    np.testing.assert_array_equal(g.numpy(), O.make_coords_grid(2, 3, 4))


def test_product_package_does_not_import_the_oracle():
    import os
    import re

    pkg = os.path.dirname(J.__file__)
    for root, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".cc", ".h")):
                text = open(os.path.join(root, f)).read()
                assert not re.search(r"^\s*(from|import)\s+oracle\b", text, flags=re.M), f


def test_checkpoint_msgpack_round_trip_and_torch_conversion(tmp_path):
    """f3: scripts/convert_checkpoint.py's path (torch state_dict -> Flax tree -> msgpack) and back."""
    import torchvision.models.optical_flow as tvof

    from jax_raft_b200 import checkpoint

    torch.manual_seed(3)
    net = tvof.raft_small(weights=None)
    pth = tmp_path / "raft_small.pth"
    torch.save(net.state_dict(), pth)
    out = tmp_path / "raft_small.msgpack"
    variables = checkpoint.convert_checkpoint(str(pth), str(out))
    loaded = checkpoint.load_msgpack(str(out))
    a, b = dict(_tree_items(variables)), dict(_tree_items(loaded))
    assert set(a) == set(b)
    for k in a:
        assert b[k].dtype == np.float32 and np.array_equal(a[k], b[k]), k
    # the tree loads back into torchvision bit-for-bit
    sd = J.torch_state_dict_from_variables(loaded)
    for k, v in net.state_dict().items():
        if not k.endswith("num_batches_tracked"):
            assert torch.equal(sd[k], v), k
    # ext-type framing as flax writes it: ExtType(1, packb((shape, dtype, bytes)))
    import msgpack

    model, from_file = J.raft_small(pretrained=str(out))  # local file in place of the download (model.py:684-689)
    assert np.array_equal(from_file["params"]["feature_encoder"]["conv"]["kernel"], a["params/feature_encoder/conv/kernel"])
    raw = msgpack.unpackb(out.read_bytes(), raw=False, strict_map_key=False)
    leaf = raw["params"]["feature_encoder"]["conv"]["bias"]
    assert isinstance(leaf, msgpack.ExtType) and leaf.code == 1
    shape, dtype, buf = msgpack.unpackb(leaf.data, raw=True)
    assert tuple(shape) == (128,) and dtype == b"float32" and len(buf) == 512


def test_validation_helpers(tmp_path):
    """f4: InputPadder / .flo / EPE metrics of scripts/validate_sintel.py."""
    from jax_raft_b200 import validation as V

    img = np.arange(436 * 1024 * 3, dtype=np.float32).reshape(1, 436, 1024, 3)
    padder = V.InputPadder(img.shape)  # Sintel 436x1024 -> 440x1024 (validate_sintel.py:23-35)
    (padded,) = padder.pad(img)
    assert padded.shape == (1, 440, 1024, 3)
    assert np.array_equal(padded[0, 0], img[0, 0]) and np.array_equal(padded[0, 1], img[0, 0])  # replicate, 2 rows on top
    assert np.array_equal(padder.unpad(padded), img)
    other = V.InputPadder((1, 370, 1226, 3), mode="kitti")
    (p2,) = other.pad(np.zeros((1, 370, 1226, 3), np.float32))
    assert p2.shape == (1, 376, 1232, 3)
    # torch's F.pad(replicate) is the reference's implementation
    ref = torch.nn.functional.pad(torch.from_numpy(img).permute(0, 3, 1, 2), [0, 0, 2, 2], mode="replicate")
    assert np.array_equal(padded, ref.permute(0, 2, 3, 1).numpy())

    flow = np.random.default_rng(0).standard_normal((7, 9, 2)).astype(np.float32)
    V.write_flo(str(tmp_path / "a.flo"), flow)
    assert np.array_equal(V.read_flo(str(tmp_path / "a.flo")), flow)
    (tmp_path / "bad.flo").write_bytes(b"\x00" * 32)
    assert V.read_flo(str(tmp_path / "bad.flo")) is None

    m = V.epe_metrics(np.array([0.5, 2.0, 4.0, 10.0]))
    assert m == {"epe": 4.125, "1px": 0.25, "3px": 0.5, "5px": 0.75}


def test_jax_ffi_layer_is_gated_not_faked():
    """The jax.ffi layer is shipped as source; without jax it must refuse to
    import rather than pretend, and its handler table must name exactly the symbols the C++ file
    defines."""
    import importlib
    import os
    import re

    try:
        importlib.import_module("jax")
        pytest.skip("jax is installed here; the gate is not exercised")
    except ImportError:
        pass
    with pytest.raises(ImportError, match="needs jax"):
        importlib.import_module("jax_raft_b200.jaxffi")
    here = os.path.join(os.path.dirname(J.__file__), "jaxffi")
    cc = open(os.path.join(here, "raftk_ffi.cc")).read()
    py = open(os.path.join(here, "__init__.py")).read()
    defined = set(re.findall(r"XLA_FFI_DEFINE_HANDLER_SYMBOL\((\w+),", cc))
    named = set(re.findall(r'"raftk_\w+": "(\w+)"', py))
    assert defined == named and len(defined) == 6
    for fn in re.findall(r"\b(raftk_\w+)\(", cc):
        assert fn in open(os.path.join(os.path.dirname(os.path.dirname(J.__file__)), "include", "raftk.h")).read(), fn


def test_weight_packing_for_the_fused_kernels():
    """Host-side layouts of the f1 / f2 kernels (pure index work, checked without a GPU):
    pack_mask_predictor: row g*144 + k*16 + t <- mask channel k*64 + 16*g + t;
    CorrBlock.pack_conv_weight: column l*KL + t <- channel l*(2r+1)^2 + t, zero padding up to KL."""
    import torch

    w = torch.arange(576 * 5, dtype=torch.float32).reshape(576, 5, 1, 1)
    b = torch.arange(576, dtype=torch.float32)
    pw, pb = J.pack_mask_predictor(w, b)
    assert pw.shape == (576, 5) and pb.shape == (576,)
    for g, k, t in ((0, 0, 0), (1, 3, 7), (3, 8, 15), (2, 5, 0)):
        ch = k * 64 + 16 * g + t
        row = g * 144 + k * 16 + t
        assert pb[row] == ch and torch.equal(pw[row], w[ch, :, 0, 0])
    assert sorted(pb.tolist()) == list(range(576))  # a permutation

    for r, kl in ((4, 96), (3, 64)):
        cb = J.CorrBlock(4, r)
        ss = (2 * r + 1) ** 2
        wc = torch.arange(32 * 4 * ss, dtype=torch.float32).reshape(32, 4 * ss, 1, 1) + 1
        packed = cb.pack_conv_weight(wc)
        assert packed.shape == (32, 4 * kl)
        for l in range(4):
            assert torch.equal(packed[:, l * kl:l * kl + ss], wc[:, l * ss:(l + 1) * ss, 0, 0])
            assert not packed[:, l * kl + ss:(l + 1) * kl].any()
    assert not J.CorrBlock(3, 4).fused_conv_supported and not J.CorrBlock(4, 2).fused_conv_supported


def test_pyramid_layout_and_storage_options():
    """The round-2 host options, without a GPU: tiled slab shapes, storage / layout / kernel argument checks, and the
    models' defaults (float16 slabs in the tf32 mode, float32 slabs in the fp32 mode, row-major float32 for f1)."""
    from jax_raft_b200.corr import TiledPyramid

    # (Q, tiles_y, tiles_x, 4, 4): tiles_y = ceil(lh / 4), tiles_x = 2 * ceil(lw / 8)
    assert TiledPyramid.level_shape(7, 55, 128, 0) == (7, 14, 32, 4, 4)
    assert TiledPyramid.level_shape(7, 55, 128, 3) == (7, 2, 4, 4, 4)
    assert TiledPyramid.level_shape(1, 46, 62, 1) == (1, 6, 8, 4, 4)   # 23 x 31
    assert TiledPyramid.level_shape(1, 136, 240, 2) == (1, 9, 16, 4, 4)  # 34 x 60
    with pytest.raises(AssertionError):
        J.CorrBlock(4, 4, storage="bf16")
    with pytest.raises(AssertionError):
        J.CorrBlock(4, 4, layout="rowmajor", storage="fp16")
    with pytest.raises(AssertionError):
        J.AltCorrBlock(4, 4, kernel="wmma")
    assert J.CorrBlock(4, 4).storage == "fp32" and J.CorrBlock(4, 4).layout == "tiled"
    assert J.raft_large()[0].corr_block.storage == "fp16"
    assert J.raft_large(precision="fp32")[0].corr_block.storage == "fp32"
    f1 = J.raft_large(fused_lookup_conv=True)[0].corr_block
    assert f1.layout == "rowmajor" and f1.storage == "fp32"
    alt = J.AltCorrBlock(4, 4)
    assert alt._tc_supported(256) and alt._tc_supported(128) and not alt._tc_supported(96) and not alt._tc_supported(512)
    assert not J.AltCorrBlock(5, 4)._tc_supported(256) and not J.AltCorrBlock(4, 2)._tc_supported(256)
    assert not J.AltCorrBlock(4, 4, kernel="simt")._tc_supported(256)
    # unsupported factory keywords of the reference (model.py:636-670) are refused, corr_block= is accepted
    with pytest.raises(TypeError):
        J.raft_large(update_block=object())
    assert J.raft_large(corr_block=alt)[0].corr_block is alt
