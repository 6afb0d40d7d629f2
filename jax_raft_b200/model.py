"""RAFT host model with the reference's Flax-shaped API, hot path on libraftk (sm_100a).

Mirrors ``/root/reference/jax_raft/model.py``:

* ``raft_large(*, pretrained=False, **kwargs) -> (model, variables)`` (``:694``), ``raft_small`` (``:732``);
  ``corr_block=`` is the injection point (``:643-644``).
* ``model.apply(variables, image1, image2, train=False, num_flow_updates=12) -> (N, B, H, W, 2)``
  (``:557``, ``:605``); images NHWC float32 in [-1, 1] (``examples/demo.py:9-10``).
* ``variables = {"params": ..., "batch_stats": ...}`` with the Flax tree names and layouts
  (conv ``kernel`` HWIO, norm ``scale``/``bias``, ``batch_stats`` ``mean``/``var``), i.e. exactly
  what ``scripts/convert_checkpoint.py:11-56`` produces from a torchvision ``state_dict``.

Scope: the correlation block and the convex upsampler are the hand-written
CUDA kernels; encoders, the conv-GRU update block and the mask predictor are dense convolutions
and deliberately stay on the host framework's library path (cuDNN through torch).  Inference only
(``train=True`` raises): there is no VJP for the kernels.
"""

from __future__ import annotations

import math
from dataclasses import dataclass, field
from typing import Any, Dict, Optional, Sequence, Tuple

import numpy as np
import torch
import torch.nn.functional as F

from . import _lib
from .corr import CorrBlock
from .functional import make_coords_grid, mask_upsample, pack_mask_predictor, upsample_flow

Tree = Dict[str, Any]


# ---------------------------------------------------------------------------------------
# Architecture description (the keyword arguments of _raft, model.py:608-635)
# ---------------------------------------------------------------------------------------
@dataclass(frozen=True)
class RaftConfig:
    name: str
    feature_encoder_layers: Tuple[int, ...]
    feature_encoder_block: str  # "residual" | "bottleneck"
    feature_encoder_norm: Optional[str]  # "instance" | "batch" | None
    context_encoder_layers: Tuple[int, ...]
    context_encoder_block: str
    context_encoder_norm: Optional[str]
    corr_block_num_levels: int
    corr_block_radius: int
    motion_encoder_corr_layers: Tuple[int, ...]
    motion_encoder_flow_layers: Tuple[int, ...]
    motion_encoder_out_channels: int
    recurrent_block_hidden_state_size: int
    recurrent_block_kernel_size: Tuple[Tuple[int, int], ...]
    recurrent_block_padding: Tuple[Tuple[int, int], ...]
    flow_head_hidden_size: int
    use_mask_predictor: bool
    encoder_strides: Tuple[int, ...] = (2, 1, 2, 2)  # FeatureEncoder.strides, model.py:227


RAFT_LARGE = RaftConfig(  # model.py:702-729
    name="raft_large",
    feature_encoder_layers=(64, 64, 96, 128, 256), feature_encoder_block="residual", feature_encoder_norm="instance",
    context_encoder_layers=(64, 64, 96, 128, 256), context_encoder_block="residual", context_encoder_norm="batch",
    corr_block_num_levels=4, corr_block_radius=4,
    motion_encoder_corr_layers=(256, 192), motion_encoder_flow_layers=(128, 64), motion_encoder_out_channels=128,
    recurrent_block_hidden_state_size=128, recurrent_block_kernel_size=((1, 5), (5, 1)),
    recurrent_block_padding=((0, 2), (2, 0)), flow_head_hidden_size=256, use_mask_predictor=True,
)

RAFT_SMALL = RaftConfig(  # model.py:740-767
    name="raft_small",
    feature_encoder_layers=(32, 32, 64, 96, 128), feature_encoder_block="bottleneck", feature_encoder_norm="instance",
    context_encoder_layers=(32, 32, 64, 96, 160), context_encoder_block="bottleneck", context_encoder_norm=None,
    corr_block_num_levels=4, corr_block_radius=3,
    motion_encoder_corr_layers=(96,), motion_encoder_flow_layers=(64, 32), motion_encoder_out_channels=82,
    recurrent_block_hidden_state_size=96, recurrent_block_kernel_size=((3, 3),),
    recurrent_block_padding=((1, 1),), flow_head_hidden_size=128, use_mask_predictor=False,
)


# ---------------------------------------------------------------------------------------
# Variable tree: shapes, initialisers, conversion to/from torchvision state_dict keys
# ---------------------------------------------------------------------------------------
def _conv_specs(cfg: RaftConfig):
    """Yield (path, kh, kw, cin, cout, init, norm) for every conv; path is the Flax tree path.

    ``init`` is "kaiming" for the reference's ``Conv()`` helper (variance_scaling(2, fan_out,
    truncated_normal), model.py:101-104) and "lecun" for bare ``nn.Conv`` (flax default).
    ``norm`` is the norm layer that follows inside a ConvNormActivation ("batch" owns params).
    """
    def encoder(prefix, layers, block, norm):
        yield (f"{prefix}/convnormrelu/layers_0", 7, 7, 3, layers[0], "kaiming", norm)
        cin = layers[0]
        for li, (cout, stride) in enumerate(zip(layers[1:4], cfg.encoder_strides[1:]), start=1):
            for bi in range(2):
                s = stride if bi == 0 else 1
                base = f"{prefix}/layer{li}/layers_{bi}"
                if block == "residual":
                    yield (f"{base}/convnormrelu1/layers_0", 3, 3, cin, cout, "kaiming", norm)
                    yield (f"{base}/convnormrelu2/layers_0", 3, 3, cout, cout, "kaiming", norm)
                else:
                    yield (f"{base}/convnormrelu1/layers_0", 1, 1, cin, cout // 4, "kaiming", norm)
                    yield (f"{base}/convnormrelu2/layers_0", 3, 3, cout // 4, cout // 4, "kaiming", norm)
                    yield (f"{base}/convnormrelu3/layers_0", 1, 1, cout // 4, cout, "kaiming", norm)
                if s != 1:
                    yield (f"{base}/downsample/layers_0", 1, 1, cin, cout, "kaiming", norm)
                cin = cout
        yield (f"{prefix}/conv", 1, 1, cin, layers[4], "kaiming", None)

    yield from encoder("feature_encoder", cfg.feature_encoder_layers, cfg.feature_encoder_block, cfg.feature_encoder_norm)
    yield from encoder("context_encoder", cfg.context_encoder_layers, cfg.context_encoder_block, cfg.context_encoder_norm)

    corr_ch = cfg.corr_block_num_levels * (2 * cfg.corr_block_radius + 1) ** 2
    me = "update_block/motion_encoder"
    cl, fl = cfg.motion_encoder_corr_layers, cfg.motion_encoder_flow_layers
    yield (f"{me}/convcorr1/layers_0", 1, 1, corr_ch, cl[0], "kaiming", None)
    if len(cl) == 2:
        yield (f"{me}/convcorr2/layers_0", 3, 3, cl[0], cl[1], "kaiming", None)
    yield (f"{me}/convflow1/layers_0", 7, 7, 2, fl[0], "kaiming", None)
    yield (f"{me}/convflow2/layers_0", 3, 3, fl[0], fl[1], "kaiming", None)
    yield (f"{me}/conv/layers_0", 3, 3, cl[-1] + fl[1], cfg.motion_encoder_out_channels - 2, "kaiming", None)

    hid = cfg.recurrent_block_hidden_state_size
    ctx = cfg.context_encoder_layers[4] - hid
    gru_in = hid + ctx + cfg.motion_encoder_out_channels
    for gi, (kh, kw) in enumerate(cfg.recurrent_block_kernel_size, start=1):
        for g in ("convz", "convr", "convq"):
            yield (f"update_block/recurrent_block/convgru{gi}/{g}", kh, kw, gru_in, hid, "lecun", None)
    yield ("update_block/flow_head/conv1", 3, 3, hid, cfg.flow_head_hidden_size, "lecun", None)
    yield ("update_block/flow_head/conv2", 3, 3, cfg.flow_head_hidden_size, 2, "lecun", None)
    if cfg.use_mask_predictor:
        yield ("mask_predictor/convrelu/layers_0", 3, 3, hid, 256, "kaiming", None)
        yield ("mask_predictor/conv", 1, 1, 256, 8 * 8 * 9, "lecun", None)


def _tree_set(tree: Tree, path: str, value) -> None:
    parts = path.split("/")
    for p in parts[:-1]:
        tree = tree.setdefault(p, {})
    tree[parts[-1]] = value


def _tree_get(tree: Tree, path: str):
    for p in path.split("/"):
        tree = tree[p]
    return tree


def _tree_items(tree: Tree, prefix: str = ""):
    for k, v in tree.items():
        path = f"{prefix}/{k}" if prefix else k
        if isinstance(v, dict):
            yield from _tree_items(v, path)
        else:
            yield path, v


def init_variables(cfg: RaftConfig, seed: int = 0) -> Tree:
    """Random-init ``{"params", "batch_stats"}`` with the reference's initialisers (host numpy).

    The reference draws from ``jax.random.PRNGKey(0)`` (model.py:681-682); JAX's generator is not
    available here, so the values differ while the distributions match.
    """
    g = torch.Generator().manual_seed(seed)
    params: Tree = {}
    stats: Tree = {}
    for path, kh, kw, cin, cout, init, norm in _conv_specs(cfg):
        fan = kh * kw * (cout if init == "kaiming" else cin)
        scale = 2.0 if init == "kaiming" else 1.0
        std = math.sqrt(scale / fan) / 0.87962566103423978
        kernel = torch.empty((kh, kw, cin, cout))
        torch.nn.init.trunc_normal_(kernel, mean=0.0, std=std, a=-2 * std, b=2 * std, generator=g)
        _tree_set(params, f"{path}/kernel", kernel.numpy())
        _tree_set(params, f"{path}/bias", np.zeros((cout,), np.float32))
        if norm == "batch":
            npath = path.rsplit("/", 1)[0] + "/layers_1"
            _tree_set(params, f"{npath}/scale", np.ones((cout,), np.float32))
            _tree_set(params, f"{npath}/bias", np.zeros((cout,), np.float32))
            _tree_set(stats, f"{npath}/mean", np.zeros((cout,), np.float32))
            _tree_set(stats, f"{npath}/var", np.ones((cout,), np.float32))
    return {"params": params, "batch_stats": stats}


def variables_from_torch_state_dict(state_dict) -> Tree:
    """torchvision RAFT ``state_dict`` -> Flax-layout variables; the rules of
    ``scripts/convert_checkpoint.py``: conv OIHW -> HWIO ``kernel`` (``:20``), 1-D ``weight`` ->
    ``scale`` (``:23``), numeric path parts ``N`` -> ``layers_N`` (``:28``), ``running_mean/var`` ->
    ``batch_stats`` ``mean/var`` (``:41-44``), ``num_batches_tracked`` dropped (``:45-46``)."""
    params: Tree = {}
    stats: Tree = {}
    for key, value in state_dict.items():
        if key.endswith(".num_batches_tracked"):
            continue
        v = np.asarray(value.detach().cpu().numpy() if hasattr(value, "detach") else value, dtype=np.float32)
        parts = [("layers_" + p) if p.isdigit() else p for p in key.split(".")]
        leaf = parts[-1]
        tree = params
        if leaf == "running_mean":
            tree, leaf = stats, "mean"
        elif leaf == "running_var":
            tree, leaf = stats, "var"
        elif leaf == "weight" and v.ndim == 4:
            leaf, v = "kernel", v.transpose(2, 3, 1, 0)
        elif leaf == "weight" and v.ndim == 1:
            leaf = "scale"
        _tree_set(tree, "/".join(parts[:-1] + [leaf]), np.ascontiguousarray(v))
    return {"params": params, "batch_stats": stats}


def torch_state_dict_from_variables(variables: Tree) -> Dict[str, torch.Tensor]:
    """Inverse of :func:`variables_from_torch_state_dict` (without ``num_batches_tracked``)."""
    out: Dict[str, torch.Tensor] = {}

    def key_of(path_parts):
        return ".".join(p[len("layers_"):] if p.startswith("layers_") else p for p in path_parts)

    for path, v in _tree_items(variables["params"]):
        parts = path.split("/")
        v = np.asarray(v)
        if parts[-1] == "kernel":
            out[key_of(parts[:-1] + ["weight"])] = torch.from_numpy(np.ascontiguousarray(v.transpose(3, 2, 0, 1)))
        elif parts[-1] == "scale":
            out[key_of(parts[:-1] + ["weight"])] = torch.from_numpy(v.copy())
        else:
            out[key_of(parts)] = torch.from_numpy(v.copy())
    for path, v in _tree_items(variables.get("batch_stats", {})):
        parts = path.split("/")
        leaf = {"mean": "running_mean", "var": "running_var"}[parts[-1]]
        out[key_of(parts[:-1] + [leaf])] = torch.from_numpy(np.asarray(v).copy())
    return out


# ---------------------------------------------------------------------------------------
# Prepared (device, library-layout) parameters
# ---------------------------------------------------------------------------------------
@dataclass
class _ConvP:
    w: torch.Tensor  # OIHW, channels_last
    b: torch.Tensor


def _prepare(cfg: RaftConfig, variables: Tree, device) -> Dict[str, _ConvP]:
    """HWIO kernels -> OIHW channels_last on ``device``; inference BatchNorm (flax epsilon 1e-5,
    use_running_average) folded into the preceding conv; MaskPredictor's 0.25 multiplier
    (model.py:400, a power of two, so exact) folded into its last conv."""
    params, stats = variables["params"], variables.get("batch_stats", {})
    out: Dict[str, _ConvP] = {}
    for path, kh, kw, cin, cout, _init, norm in _conv_specs(cfg):
        k = torch.as_tensor(np.asarray(_tree_get(params, f"{path}/kernel")), dtype=torch.float32)
        assert tuple(k.shape) == (kh, kw, cin, cout), f"{path}/kernel: expected {(kh, kw, cin, cout)}, got {tuple(k.shape)}"
        b = torch.as_tensor(np.asarray(_tree_get(params, f"{path}/bias")), dtype=torch.float32)
        w = k.permute(3, 2, 0, 1).contiguous().to(torch.float64)
        b = b.to(torch.float64)
        if norm == "batch":
            npath = path.rsplit("/", 1)[0] + "/layers_1"
            scale = torch.as_tensor(np.asarray(_tree_get(params, f"{npath}/scale")), dtype=torch.float64)
            nbias = torch.as_tensor(np.asarray(_tree_get(params, f"{npath}/bias")), dtype=torch.float64)
            mean = torch.as_tensor(np.asarray(_tree_get(stats, f"{npath}/mean")), dtype=torch.float64)
            var = torch.as_tensor(np.asarray(_tree_get(stats, f"{npath}/var")), dtype=torch.float64)
            g = scale / torch.sqrt(var + 1e-5)
            w = w * g[:, None, None, None]
            b = (b - mean) * g + nbias
        if path == "mask_predictor/conv":
            w, b = w * 0.25, b * 0.25
        out[path] = _ConvP(
            w.to(torch.float32).to(device).contiguous(memory_format=torch.channels_last),
            b.to(torch.float32).to(device),
        )
    # Channel counts the library would pad with kernels of its own (nhwcAddPaddingKernel x2 + memset per
    # call: 126 or 2 output channels; a non-tensor-core engine for 2 input channels) are padded here
    # once, with zero filters, to the next multiple of 8 (profiles/conv_kernels.py: 21 -> 11 us, 16.5 -> 11 us,
    # 16 -> 9.5 us per call at 55x128)
    def pad8(n):
        return (n + 7) // 8 * 8

    def padded(p, cout=None, cin=None):
        w, b = p.w, p.b
        if cout is not None and cout > w.shape[0]:
            w = torch.cat([w, w.new_zeros((cout - w.shape[0],) + tuple(w.shape[1:]))], dim=0)
            b = torch.cat([b, b.new_zeros(cout - b.shape[0])])
        if cin is not None and cin > w.shape[1]:
            w = torch.cat([w, w.new_zeros((w.shape[0], cin - w.shape[1]) + tuple(w.shape[2:]))], dim=1)
        return _ConvP(w.contiguous(memory_format=torch.channels_last), b.contiguous())

    me = "update_block/motion_encoder"
    out[f"{me}/conv/layers_0/padded"] = padded(out[f"{me}/conv/layers_0"], cout=pad8(cfg.motion_encoder_out_channels - 2))
    out["update_block/flow_head/conv2/padded"] = padded(out["update_block/flow_head/conv2"], cout=8)
    out[f"{me}/convflow1/layers_0/padded"] = padded(out[f"{me}/convflow1/layers_0"], cin=8)
    # ConvGRU: the z and r filters read the same input (model.py:306-307); stacked, they are one convolution
    for gi in range(1, len(cfg.recurrent_block_kernel_size) + 1):
        base = f"update_block/recurrent_block/convgru{gi}"
        z, r = out[f"{base}/convz"], out[f"{base}/convr"]
        out[f"{base}/convzr"] = _ConvP(torch.cat([z.w, r.w], dim=0).contiguous(memory_format=torch.channels_last),
                                       torch.cat([z.b, r.b], dim=0).contiguous())
    return out


# ---------------------------------------------------------------------------------------
# Forward pieces (NCHW-shaped channels_last tensors == NHWC memory)
# ---------------------------------------------------------------------------------------
_ACT = {None: 0, "relu": 1, "sigmoid": 2, "tanh": 3}
# conv + bias + relu as the library's fused convolution (one launch instead of conv + bias_act);
# RAFTK_CUDNN_FUSED_RELU=0 restores the two-launch form for A/B timing
_CUDNN_FUSED_RELU = bool(int(__import__("os").environ.get("RAFTK_CUDNN_FUSED_RELU", "1")))


def _conv(x, p: _ConvP, stride=1, padding=0, act=None, bias=True):
    """Library convolution + (bias, activation) epilogue.  The bias is not given to the convolution:
    on channels_last tensors torch adds it with a strided broadcast kernel that costs more than the
    1x1 convolutions themselves; ``raftk_bias_act_nhwc`` applies bias and activation in one pass."""
    if _CUDNN_FUSED_RELU and act == "relu" and bias:
        pad = (padding, padding) if isinstance(padding, int) else tuple(padding)
        st = (stride, stride) if isinstance(stride, int) else tuple(stride)
        return torch.cudnn_convolution_relu(x, p.w, p.b, st, pad, (1, 1), 1)
    y = F.conv2d(x, p.w, None, stride=stride, padding=padding)
    if not bias and act is None:
        return y
    if not y.is_contiguous(memory_format=torch.channels_last):
        y = y.contiguous(memory_format=torch.channels_last)
    n, c, h, w = y.shape
    if not bias:
        zero = getattr(p, "_zero_bias", None)
        if zero is None:
            zero = torch.zeros_like(p.b)
            p._zero_bias = zero
        b = zero
    else:
        b = p.b
    lib = _lib.load()
    with torch.cuda.device(y.device):
        rc = lib.raftk_bias_act_nhwc(y.data_ptr(), b.data_ptr(), n * h * w, c, _ACT[act], _lib.current_stream_ptr(y.device))
    _lib.check("raftk_bias_act_nhwc", rc)
    _lib.count_launch(1)
    return y


def _instance_norm_(y, relu: bool, eps: float = 1e-5, residual=None):
    """In-place InstanceNorm(+ReLU) of a conv output through the glue kernel; with ``residual`` also the
    residual block's tail ``relu(residual + .)``.  Shapes it does not cover (channel counts not a
    multiple of 4) use the library ops."""
    n, c, h, w = y.shape
    if c % 4 or c > 1024:
        y = F.instance_norm(y, eps=eps)
        y = F.relu(y) if relu else y
        return y if residual is None else F.relu(residual + y)
    y = _cl(y)
    lib = _lib.load()
    nbytes = lib.raftk_instance_norm_scratch_bytes(n, c)
    scratch = torch.empty((nbytes,), dtype=torch.uint8, device=y.device)
    with torch.cuda.device(y.device):
        if residual is None:
            rc = lib.raftk_instance_norm_nhwc(y.data_ptr(), n, h * w, c, eps, 1 if relu else 0, scratch.data_ptr(), nbytes,
                                              _lib.current_stream_ptr(y.device))
        else:
            residual = _cl(residual)
            assert residual.shape == y.shape
            rc = lib.raftk_instance_norm_residual_nhwc(y.data_ptr(), residual.data_ptr(), n, h * w, c, eps, 1 if relu else 0,
                                                       scratch.data_ptr(), nbytes, _lib.current_stream_ptr(y.device))
    _lib.check("raftk_instance_norm_nhwc", rc)
    _lib.count_launch(2)
    return y


def _add_relu_(x, y):
    """relu(x + y) in one pass, into ``x`` when it is a dense channels_last tensor this call may overwrite."""
    if x.numel() % 4 or not x.is_contiguous(memory_format=torch.channels_last) or not y.is_contiguous(memory_format=torch.channels_last):
        return F.relu(x + y)
    lib = _lib.load()
    with torch.cuda.device(x.device):
        rc = lib.raftk_add_relu(x.data_ptr(), y.data_ptr(), x.numel(), _lib.current_stream_ptr(x.device))
    _lib.check("raftk_add_relu", rc)
    _lib.count_launch(1)
    return x


def _cna(P, path, x, norm, stride=1, padding=None, act=True, residual=None):
    """ConvNormActivation (model.py:120-159): conv (+bias), norm, relu; with ``residual`` (instance norm
    only) also the enclosing block's ``relu(residual + .)``."""
    p = P[f"{path}/layers_0"]
    if padding is None:
        padding = ((p.w.shape[2] - 1) // 2, (p.w.shape[3] - 1) // 2)
    if norm == "instance":  # nn.InstanceNorm(epsilon=1e-5, no scale/bias), model.py:706-707
        # a per-channel bias cancels in (x - mean) / std, so the conv bias is not applied at all
        return _instance_norm_(_conv(x, p, stride, padding, bias=False), relu=act, residual=residual)
    assert residual is None
    return _conv(x, p, stride, padding, act="relu" if act else None)


def _encoder(P, prefix, x, layers, block, norm, strides):
    x = _cna(P, f"{prefix}/convnormrelu", x, norm, stride=strides[0])
    for li, stride in enumerate(strides[1:], start=1):
        for bi in range(2):
            s = stride if bi == 0 else 1
            base = f"{prefix}/layer{li}/layers_{bi}"
            last = "convnormrelu2" if block == "residual" else "convnormrelu3"
            if s != 1:  # the shortcut first, so that the block's last norm can absorb the residual tail
                shortcut = _cna(P, f"{base}/downsample", x, norm, stride=s, act=False)
            else:
                shortcut = x
            fuse_tail = norm == "instance"
            if block == "residual":  # model.py:162-184
                y = _cna(P, f"{base}/convnormrelu1", x, norm, stride=s)
            else:  # model.py:187-216
                y = _cna(P, f"{base}/convnormrelu1", x, norm)
                y = _cna(P, f"{base}/convnormrelu2", y, norm, stride=s)
            y = _cna(P, f"{base}/{last}", y, norm, residual=shortcut if fuse_tail else None)
            # relu(shortcut + y): inside the norm kernel above, else one pass into y (a fresh conv output)
            x = y if fuse_tail else _add_relu_(y, shortcut)
    return _conv(x, P[f"{prefix}/conv"])


def _motion_encoder(P, cfg, flow, corr_features):
    """model.py:260-290."""
    me = "update_block/motion_encoder"
    corr = _cna(P, f"{me}/convcorr1", corr_features, None)
    if len(cfg.motion_encoder_corr_layers) == 2:
        corr = _cna(P, f"{me}/convcorr2", corr, None)
    fl = _cna(P, f"{me}/convflow1", flow, None)
    fl = _cna(P, f"{me}/convflow2", fl, None)
    cf = _cna(P, f"{me}/conv", torch.cat([corr, fl], dim=1), None)
    return torch.cat([cf, flow], dim=1)


def _cl(y):
    return y if y.is_contiguous(memory_format=torch.channels_last) else y.contiguous(memory_format=torch.channels_last)


class _RecurrentState:
    """The two ConvGRU convolution input buffers with the hidden state, context and motion features in
    channel slices (csrc/gru_glue.cu), so UpdateCell's three concatenations (model.py:289, 305, 308, 501)
    are never materialised.  ``A = [h | context | motion | flow]``, ``B = [r*h | context | motion | flow]``."""

    def __init__(self, cfg: RaftConfig, hidden: torch.Tensor, context: torch.Tensor):
        n, hid, h, w = hidden.shape
        ctx = context.shape[1]
        self.hid, self.ctx, self.c_cf = hid, ctx, cfg.motion_encoder_out_channels - 2
        self.ld = hid + ctx + cfg.motion_encoder_out_channels
        self.npix = n * h * w
        opts = dict(dtype=torch.float32, device=hidden.device, memory_format=torch.channels_last)
        self.A = torch.empty((n, self.ld, h, w), **opts)
        self.B = torch.empty((n, self.ld, h, w), **opts)
        self.z = torch.empty((n, hid, h, w), **opts)
        self.h = torch.empty((n, hid, h, w), **opts)  # dense copy of the hidden state for the heads
        self.flow = torch.zeros((n, h, w, 2), dtype=torch.float32, device=hidden.device)  # coords1 - coords0
        # the same flow in channels 0-1 of an 8-channel tensor: input of the zero-padded 7x7 flow convolution
        self.flow8 = torch.empty((n, 8, h, w), **opts).zero_()
        self.pending_delta = None  # (flow-head output, its parameters) of an iteration whose update is still to run
        self.A[:, :hid] = hidden
        self.h.copy_(hidden)
        self.A[:, hid:hid + ctx] = context
        self.B[:, hid:hid + ctx] = context

    @staticmethod
    def supported(cfg: RaftConfig) -> bool:
        hid = cfg.recurrent_block_hidden_state_size
        ld = cfg.context_encoder_layers[-1] + cfg.motion_encoder_out_channels  # hid + ctx + motion
        return hid % 2 == 0 and ld % 2 == 0 and cfg.motion_encoder_out_channels % 2 == 0

    def scatter_motion(self, cf_pre, p, flow_nhwc):
        lib = _lib.load()
        off = 4 * (self.hid + self.ctx)
        with torch.cuda.device(cf_pre.device):
            rc = lib.raftk_motion_scatter(cf_pre.data_ptr(), cf_pre.shape[1], p.b.data_ptr(), flow_nhwc.data_ptr(),
                                          self.A.data_ptr() + off, self.B.data_ptr() + off, self.ld, self.npix, self.c_cf,
                                          _lib.current_stream_ptr(cf_pre.device))
        _lib.check("raftk_motion_scatter", rc)
        _lib.count_launch(1)

    def update_flow(self, delta_pre, p, coords0, coords1):
        """model.py:504-506: coords1 += conv + bias (in place); self.flow = coords1 - coords0."""
        lib = _lib.load()
        with torch.cuda.device(coords1.device):
            rc = lib.raftk_flow_update(delta_pre.data_ptr(), delta_pre.shape[1], p.b.data_ptr(), coords0.data_ptr(),
                                       coords1.data_ptr(), self.flow.data_ptr(), self.flow8.data_ptr(), 8, self.npix,
                                       _lib.current_stream_ptr(coords1.device))
        _lib.check("raftk_flow_update", rc)
        _lib.count_launch(1)

    def gru(self, P, path, padding):
        """model.py:293-312 on the slices."""
        lib = _lib.load()
        pzr, pq = P[f"{path}/convzr"], P[f"{path}/convq"]
        dev = self.A.device
        zr_pre = _cl(F.conv2d(self.A, pzr.w, None, padding=padding))
        with torch.cuda.device(dev):
            rc = lib.raftk_gru_gate(zr_pre.data_ptr(), pzr.b.data_ptr(), self.A.data_ptr(), self.ld, self.z.data_ptr(),
                                    self.B.data_ptr(), self.ld, self.npix, self.hid, _lib.current_stream_ptr(dev))
        _lib.check("raftk_gru_gate", rc)
        q_pre = _cl(F.conv2d(self.B, pq.w, None, padding=padding))
        with torch.cuda.device(dev):
            rc = lib.raftk_gru_update(q_pre.data_ptr(), pq.b.data_ptr(), self.z.data_ptr(), self.A.data_ptr(), self.ld,
                                      self.h.data_ptr(), self.npix, self.hid, _lib.current_stream_ptr(dev))
        _lib.check("raftk_gru_update", rc)
        _lib.count_launch(2)


def _conv_gru(P, path, h, x, padding):
    """model.py:293-312."""
    hx = torch.cat([h, x], dim=1)
    z = _conv(hx, P[f"{path}/convz"], padding=padding, act="sigmoid")
    r = _conv(hx, P[f"{path}/convr"], padding=padding, act="sigmoid")
    q = _conv(torch.cat([r * h, x], dim=1), P[f"{path}/convq"], padding=padding, act="tanh")
    return torch.lerp(h, q, z)  # (1 - z) * h + z * q (model.py:311) in one pass: h + z * (q - h)


class RAFT:
    """Counterpart of the reference's ``RAFT`` module (model.py:513-605) + ``UpdateCell`` (484-510)."""

    def __init__(self, config: RaftConfig, corr_block=None, precision: str = "tf32", fused_update: bool = True,
                 concurrent: bool = True, fused_lookup_conv: bool = False, fused_mask_upsample: bool = True,
                 fused_flow_update: bool = True):
        assert precision in ("tf32", "fp32")
        # fused_update: UpdateCell's element-wise work through csrc/gru_glue.cu (same arithmetic as the
        # unfused path, which stays for configurations whose channel counts are odd and for A/B tests)
        self.fused_update = fused_update and _RecurrentState.supported(config)
        # concurrent: independent branches (context || feature encoder, flow || correlation branch of the
        # motion encoder, mask predictor + upsampler || flow head and the next iteration) are issued on
        # side streams; captured into the CUDA graph they become parallel branches.  At batch 1 no single
        # convolution fills 148 SMs, so the branches overlap almost for free.
        self.concurrent = concurrent and self.fused_update
        self._streams: Dict[str, Any] = {}
        self.config = config
        # the pyramid is opaque to the model (model.py:530-539): tiled slabs, unless the optional lookup+convolution
        # kernel (which reads row-major levels) is asked for.  In the throughput mode (precision="tf32") the slabs
        # hold float16 (each correlation value rounded once, <= 4.9e-4 relative against the north-star's 1e-3:
        # ~1e-3 px of end-to-end EPE where the TF32 convolutions contribute ~9e-3); precision="fp32" keeps
        # float32 slabs and reproduces the CPU reference to ~2e-5 px.
        self.corr_block = corr_block or CorrBlock(
            config.corr_block_num_levels, config.corr_block_radius, layout="rowmajor" if fused_lookup_conv else "tiled",
            storage="fp16" if precision == "tf32" and not fused_lookup_conv else "fp32")
        expected = config.corr_block_num_levels * (2 * config.corr_block_radius + 1) ** 2
        assert self.corr_block.out_channels == expected, (
            f"corr_block.out_channels={self.corr_block.out_channels} does not match the motion encoder ({expected})")
        # "tf32": cuDNN may use TF32 for the float32 convolutions — what XLA's GPU backend does
        # by default for the reference.  "fp32": full-precision library convolutions.
        self.precision = precision
        # fused_lookup_conv: the lookup and MotionEncoder.convcorr1 as one kernel (raftk_corr_lookup_conv1x1,
        # SURVEY §8 f1).  Off by default: measured slower than the two kernels it replaces.
        # fused_flow_update: UpdateCell's tail (coords1 += delta; flow = coords1 - coords0) inside the next
        # iteration's lookup kernel (raftk_corr_lookup_update) instead of a launch of its own
        self.fused_flow_update = (fused_flow_update and self.concurrent and not fused_lookup_conv
                                  and hasattr(self.corr_block, "index_pyramid_update")
                                  and getattr(self.corr_block, "fused_conv_supported", False))
        # fused_mask_upsample: MaskPredictor's last 1x1 convolution inside the upsampler (raftk_mask_upsample,
        # SURVEY §8 f2): parity at batch 1, 40 % faster than convolution + upsampler from batch 4 up
        hidden_mask = 256  # MaskPredictor's hidden size (model.py:383)
        self.fused_mask_upsample = (fused_mask_upsample and self.fused_update and precision == "tf32"
                                    and config.use_mask_predictor and hidden_mask % 32 == 0 and hidden_mask <= 256)
        self.fused_lookup_conv = (fused_lookup_conv and self.fused_update and precision == "tf32"
                                  and getattr(self.corr_block, "fused_conv_supported", False)
                                  and config.motion_encoder_corr_layers[0] % 16 == 0 and config.motion_encoder_corr_layers[0] <= 256)
        self._prepared: Dict[Tuple[int, str], Any] = {}
        self._graphs: Dict[Any, "_GraphRunner"] = {}

    # Flax-style init: returns fresh variables (model.py:681-682)
    def init(self, seed: int = 0) -> Tree:
        return init_variables(self.config, seed)

    def invalidate(self) -> None:
        """Forget prepared weights and captured graphs.  ``prepare`` and the graph cache key on the identity
        of the ``variables`` tree, so call this after changing a tree's arrays IN PLACE (a new tree object
        needs nothing)."""
        self._prepared.clear()
        self._graphs.clear()

    def prepare(self, variables: Tree, device) -> Dict[str, _ConvP]:
        key = (id(variables), str(device))
        if key not in self._prepared:
            # hold a reference to `variables` so its id cannot be reused while cached
            self._prepared[key] = (variables, _prepare(self.config, variables, device))
        return self._prepared[key][1]

    # -- model.py:557-605 ---------------------------------------------------------------
    def forward(self, P, image1: torch.Tensor, image2: torch.Tensor, num_flow_updates: int) -> torch.Tensor:
        cfg = self.config
        batch_size, h, w, _ = image1.shape
        assert (h, w) == tuple(image2.shape[-3:-1]), "input images should have the same shape"
        assert (h % 8 == 0) and (w % 8 == 0), "input image H and W should be divisible by 8"
        prev = torch.backends.cudnn.allow_tf32
        prev_bench = torch.backends.cudnn.benchmark
        torch.backends.cudnn.allow_tf32 = self.precision == "tf32"
        torch.backends.cudnn.benchmark = True  # the library times its engines once per convolution shape
        try:
            main = torch.cuda.current_stream(image1.device)
            sa, sb = self._side_streams(image1.device) if self.concurrent else (None, None)
            hid = cfg.recurrent_block_hidden_state_size

            def context_branch():
                context_out = _encoder(P, "context_encoder", image1.permute(0, 3, 1, 2), cfg.context_encoder_layers,
                                       cfg.context_encoder_block, cfg.context_encoder_norm, cfg.encoder_strides)
                assert context_out.shape[1] - hid > 0
                return torch.tanh(context_out[:, :hid]), F.relu(context_out[:, hid:])

            if sa is not None:
                sa.wait_stream(main)
                with torch.cuda.stream(sa):
                    hidden_state, context = context_branch()

            x = torch.cat([image1, image2], dim=0).permute(0, 3, 1, 2)  # NHWC memory, NCHW view
            fmaps = _encoder(P, "feature_encoder", x, cfg.feature_encoder_layers, cfg.feature_encoder_block,
                             cfg.feature_encoder_norm, cfg.encoder_strides)
            fmaps = fmaps.permute(0, 2, 3, 1).contiguous()  # (2B, h/8, w/8, D) NHWC
            fmap1, fmap2 = fmaps[:batch_size], fmaps[batch_size:]
            assert tuple(fmap1.shape[-3:-1]) == (h // 8, w // 8), "The feature encoder should downsample H and W by 8"

            corr_pyramid = self.corr_block.build_pyramid(fmap1, fmap2)  # model.py:567

            if sa is not None:
                main.wait_stream(sa)
            else:
                hidden_state, context = context_branch()

            coords0 = make_coords_grid(batch_size, h // 8, w // 8, device=image1.device)
            coords1 = coords0.clone()
            state = _RecurrentState(cfg, hidden_state, context) if self.fused_update else None

            flows, out = [], None
            if state is not None and cfg.use_mask_predictor:  # the upsampler writes the stacked result directly
                out = torch.empty((num_flow_updates, batch_size, h, w, 2), dtype=torch.float32, device=image1.device)
            pending = None  # the previous iteration's upsampler launch, deferred past this iteration's lookup
            for i in range(num_flow_updates):  # nn.scan(UpdateCell), model.py:589-603
                if state is not None:
                    up, pending = self._update_cell_fused(P, corr_pyramid, coords0, coords1, state,
                                                          None if out is None else out[i], main, sa, sb, i == 0, pending)
                else:
                    coords1, hidden_state, up = self._update_cell(P, corr_pyramid, coords0, coords1, hidden_state, context)
                flows.append(up)
            if state is not None and state.pending_delta is not None:  # the last iteration's coordinate update
                state.update_flow(*state.pending_delta, coords0, coords1)
                state.pending_delta = None
            if pending is not None:
                pending()
            if sa is not None:  # the last mask predictor + upsampler
                main.wait_stream(sa)
            return out if out is not None else torch.stack(flows, dim=0)
        finally:
            torch.backends.cudnn.allow_tf32 = prev
            torch.backends.cudnn.benchmark = prev_bench

    def _side_streams(self, device):
        key = str(device)
        if key not in self._streams:
            self._streams[key] = (torch.cuda.Stream(device=device), torch.cuda.Stream(device=device))
        return self._streams[key]

    # -- model.py:494-510 ---------------------------------------------------------------
    def _update_cell(self, P, corr_pyramid, coords0, coords1, hidden_state, context):
        """UpdateCell as the reference composes it (concatenations, one op per line)."""
        cfg = self.config
        corr_features = self.corr_block.index_pyramid(corr_pyramid, coords1)  # (B, h, w, C) NHWC
        flow = (coords1 - coords0).permute(0, 3, 1, 2)
        motion = _motion_encoder(P, cfg, flow, corr_features.permute(0, 3, 1, 2))
        x = torch.cat([context, motion], dim=1)
        for gi, pad in enumerate(cfg.recurrent_block_padding, start=1):
            hidden_state = _conv_gru(P, f"update_block/recurrent_block/convgru{gi}", hidden_state, x, pad)
        d = _conv(hidden_state, P["update_block/flow_head/conv1"], padding=1, act="relu")
        delta_flow = _conv(d, P["update_block/flow_head/conv2"], padding=1)
        coords1 = coords1 + delta_flow.permute(0, 2, 3, 1)
        up_mask = None
        if cfg.use_mask_predictor:  # model.py:377-400
            m = _cna(P, "mask_predictor/convrelu", hidden_state, None)
            up_mask = _conv(m, P["mask_predictor/conv"]).permute(0, 2, 3, 1).contiguous()
        upsampled = upsample_flow(flow=(coords1 - coords0).contiguous(), up_mask=up_mask)
        return coords1, hidden_state, upsampled

    def _update_cell_fused(self, P, corr_pyramid, coords0, coords1, state, out, main, sa, sb, first, pending=None):
        """The same cell on the channel-slice buffers (``_RecurrentState``); coords1 and state are
        updated in place.  With side streams ``sa`` / ``sb``: the flow branch of the motion encoder runs
        beside the correlation branch, the mask predictor beside the flow head, and the upsampler beside
        the NEXT iteration's motion-encoder convolutions: its launch is returned as ``pending`` and issued
        by the next call right after the lookup, so the two HBM-bound kernels never share the machine
        (joined before the GRU overwrites the hidden state and the flow).  Returns (upsampled, pending)."""
        cfg = self.config
        me = "update_block/motion_encoder"

        def flow_branch():
            pf = P[f"{me}/convflow1/layers_0/padded"]
            pad = ((pf.w.shape[2] - 1) // 2, (pf.w.shape[3] - 1) // 2)
            return _cna(P, f"{me}/convflow2", _conv(state.flow8, pf, padding=pad, act="relu"), None)

        fuse = self.fused_lookup_conv
        if fuse:  # lookup + convcorr1 + ReLU in one kernel (TF32 product)
            pc = P[f"{me}/convcorr1/layers_0"]
            packed = getattr(pc, "_packed", None)
            if packed is None:
                packed = pc._packed = self.corr_block.pack_conv_weight(pc.w)
            corr = self.corr_block.index_pyramid_conv(corr_pyramid, coords1, packed, pc.b, relu=True).permute(0, 3, 1, 2)
        elif state.pending_delta is not None:  # the previous iteration's coordinate update rides on this lookup
            dp, pd = state.pending_delta
            corr_features = self.corr_block.index_pyramid_update(corr_pyramid, coords0, coords1, dp, pd.b, state.flow,
                                                                 state.flow8)
            state.pending_delta = None
        else:
            corr_features = self.corr_block.index_pyramid(corr_pyramid, coords1)  # (B, h, w, C) NHWC
        if sb is not None:  # forked after the lookup, which is HBM-bound and wants the machine to itself
            sb.wait_stream(main)
            with torch.cuda.stream(sb):
                fl = flow_branch()
        if pending is not None:
            pending()
        if not fuse:
            corr = _cna(P, f"{me}/convcorr1", corr_features.permute(0, 3, 1, 2), None)
        if len(cfg.motion_encoder_corr_layers) == 2:
            corr = _cna(P, f"{me}/convcorr2", corr, None)
        if sb is not None:
            main.wait_stream(sb)
        else:
            fl = flow_branch()
        p = P[f"{me}/conv/layers_0/padded"]
        pad = ((p.w.shape[2] - 1) // 2, (p.w.shape[3] - 1) // 2)
        cf_pre = _cl(F.conv2d(torch.cat([corr, fl], dim=1), p.w, None, padding=pad))
        state.scatter_motion(cf_pre, p, state.flow)
        if sa is not None and not first:
            main.wait_stream(sa)  # the previous iteration's mask predictor / upsampler still read state.h, state.flow
        for gi, gpad in enumerate(cfg.recurrent_block_padding, start=1):
            state.gru(P, f"update_block/recurrent_block/convgru{gi}", gpad)

        fuse_mask = self.fused_mask_upsample
        if fuse_mask:  # the last 1x1 convolution inside the upsampler (TF32 product)
            pm = P["mask_predictor/conv"]
            packed = getattr(pm, "_packed", None)
            if packed is None:
                packed = pm._packed = pack_mask_predictor(pm.w, pm.b)

        def mask_branch():  # model.py:377-400; the upsampler adds the last bias itself
            m = _cna(P, "mask_predictor/convrelu", state.h, None)
            if fuse_mask:
                return m.permute(0, 2, 3, 1)  # the hidden activation: the fused kernel applies the last convolution
            return _cl(F.conv2d(m, P["mask_predictor/conv"].w, None)).permute(0, 2, 3, 1)

        def upsample(mask_or_features):
            if fuse_mask:
                return mask_upsample(state.flow, mask_or_features, packed[0], packed[1], out=out)
            return upsample_flow(state.flow, mask_or_features, mask_bias=P["mask_predictor/conv"].b, out=out)

        up_mask = None
        if cfg.use_mask_predictor and sa is not None:
            sa.wait_stream(main)
            with torch.cuda.stream(sa):
                up_mask = mask_branch()
        d = _conv(state.h, P["update_block/flow_head/conv1"], padding=1, act="relu")
        p = P["update_block/flow_head/conv2/padded"]
        delta_pre = _cl(F.conv2d(d, p.w, None, padding=1))
        if self.fused_flow_update and sa is not None and cfg.use_mask_predictor:
            # coords1 += delta, flow = coords1 - coords0 happen inside the NEXT iteration's lookup kernel
            # (raftk_corr_lookup_update); everything that reads the new flow is issued after that lookup
            state.pending_delta = (delta_pre, p)
        else:
            state.update_flow(delta_pre, p, coords0, coords1)
        if not cfg.use_mask_predictor:
            return upsample_flow(flow=state.flow), None
        if sa is None:
            return upsample(mask_branch()), None

        def deferred_upsample():
            sa.wait_stream(main)
            with torch.cuda.stream(sa):
                upsample(up_mask)

        return out, deferred_upsample

    # -- public entry, mirrors flax Module.apply -------------------------------------------
    def apply(self, variables: Tree, image1, image2, train: bool = False, num_flow_updates: int = 12,
              device=None, use_cuda_graph: bool = True) -> torch.Tensor:
        """Returns the ``num_flow_updates`` upsampled flows, shape (N, B, H, W, 2), on the GPU.

        ``image1`` / ``image2``: NHWC float32 in [-1, 1]; torch CUDA tensors, or host arrays
        (numpy / CPU tensors), which are copied to ``device`` (default: current CUDA device).
        With ``use_cuda_graph`` the whole forward is captured once per (shape, N) and replayed.
        """
        if train:
            raise NotImplementedError("inference only: the sm_100a kernels define no VJP (train=True unsupported)")
        dev = torch.device(device) if device is not None else None
        img1 = _to_device(image1, dev)
        img2 = _to_device(image2, img1.device)
        assert img1.ndim == 4 and img1.shape[-1] == 3, "images must be (B, H, W, 3)"
        P = self.prepare(variables, img1.device)
        if not use_cuda_graph:
            with torch.no_grad():
                return self.forward(P, img1, img2, num_flow_updates)
        key = (id(variables), str(img1.device), tuple(img1.shape), num_flow_updates)
        runner = self._graphs.get(key)
        if runner is None:
            runner = _GraphRunner(self, P, tuple(img1.shape), img1.device, num_flow_updates)
            self._graphs[key] = runner
        return runner(img1, img2).clone()

    def runner(self, variables: Tree, batch_size: int, height: int, width: int, num_flow_updates: int,
               device=None) -> "_GraphRunner":
        """The captured forward for one shape, with static input/output buffers (no clone)."""
        dev = torch.device(device) if device is not None else torch.device("cuda", torch.cuda.current_device())
        P = self.prepare(variables, dev)
        shape = (batch_size, height, width, 3)
        key = (id(variables), str(dev), shape, num_flow_updates)
        if key not in self._graphs:
            self._graphs[key] = _GraphRunner(self, P, shape, dev, num_flow_updates)
        return self._graphs[key]


def _to_device(x, device) -> torch.Tensor:
    if isinstance(x, torch.Tensor):
        t = x
    else:
        t = torch.from_numpy(np.ascontiguousarray(np.asarray(x, dtype=np.float32)))
    if t.dtype != torch.float32:
        t = t.to(torch.float32)
    if t.is_cuda and (device is None or t.device == device):
        return t.contiguous()
    if device is None:
        if not torch.cuda.is_available():
            raise RuntimeError("jax_raft_b200 needs a CUDA device (sm_100a); there is no CPU fallback")
        device = torch.device("cuda", torch.cuda.current_device())
    return t.to(device, non_blocking=True).contiguous()


class _GraphRunner:
    """One CUDA graph of the whole forward for a fixed (B, H, W, N): static buffers, replay."""

    def __init__(self, model: RAFT, P, shape, device, num_flow_updates: int):
        self.device = device
        # the captured kernels read the prepared weights and use the model's side streams: keep both alive
        # for as long as the graph can be replayed, even if the caller drops the model
        self._model, self._params = model, P
        self.image1 = torch.zeros(shape, dtype=torch.float32, device=device)
        self.image2 = torch.zeros(shape, dtype=torch.float32, device=device)
        with torch.cuda.device(device), torch.no_grad():
            side = torch.cuda.Stream(device=device)
            side.wait_stream(torch.cuda.current_stream(device))
            with torch.cuda.stream(side):
                for _ in range(2):  # warm-up: cuDNN autotune, lazy module loading
                    model.forward(P, self.image1, self.image2, num_flow_updates)
            torch.cuda.current_stream(device).wait_stream(side)
            torch.cuda.synchronize(device)
            self.graph = torch.cuda.CUDAGraph()
            with torch.cuda.graph(self.graph):
                self.flows = model.forward(P, self.image1, self.image2, num_flow_updates)

    def __call__(self, image1: torch.Tensor, image2: torch.Tensor) -> torch.Tensor:
        self.image1.copy_(image1, non_blocking=True)
        self.image2.copy_(image2, non_blocking=True)
        self.graph.replay()
        return self.flows


# ---------------------------------------------------------------------------------------
# Factories (model.py:608-767)
# ---------------------------------------------------------------------------------------
def _raft(cfg: RaftConfig, *, pretrained: bool = False, seed: int = 0, **kwargs):
    corr_block = kwargs.pop("corr_block", None)
    precision = kwargs.pop("precision", "tf32")
    fused_update = kwargs.pop("fused_update", True)
    concurrent = kwargs.pop("concurrent", True)
    fused_lookup_conv = kwargs.pop("fused_lookup_conv", False)
    fused_mask_upsample = kwargs.pop("fused_mask_upsample", True)
    fused_flow_update = kwargs.pop("fused_flow_update", True)
    if kwargs:
        raise TypeError(
            f"unsupported keyword(s) {sorted(kwargs)}: this build replaces the correlation block "
            "(corr_block=) only; encoders, update block and mask predictor are fixed by the architecture")
    if isinstance(pretrained, (str, bytes)) or hasattr(pretrained, "__fspath__"):
        # a local copy of the reference's released .msgpack (model.py:17-21) or of a converted one
        from .checkpoint import load_msgpack

        model = RAFT(cfg, corr_block=corr_block, precision=precision, fused_update=fused_update, concurrent=concurrent,
                     fused_lookup_conv=fused_lookup_conv, fused_mask_upsample=fused_mask_upsample, fused_flow_update=fused_flow_update)
        return model, load_msgpack(pretrained)
    if pretrained:
        # model.py:684-689 downloads a msgpack from GitHub; this environment has no network.
        raise RuntimeError(
            "pretrained=True needs network access (model.py:17-21, 684-689); pass pretrained='<file>.msgpack' "
            "(a local copy of the released checkpoint) or use variables_from_torch_state_dict(torch.load(path))")
    model = RAFT(cfg, corr_block=corr_block, precision=precision, fused_update=fused_update, concurrent=concurrent,
                     fused_lookup_conv=fused_lookup_conv, fused_mask_upsample=fused_mask_upsample, fused_flow_update=fused_flow_update)
    return model, model.init(seed)


def raft_large(*, pretrained: bool = False, **kwargs):
    """RAFT (large).  Returns ``(model, variables)`` like model.py:694-729."""
    return _raft(RAFT_LARGE, pretrained=pretrained, **kwargs)


def raft_small(*, pretrained: bool = False, **kwargs):
    """RAFT (small).  Returns ``(model, variables)`` like model.py:732-767."""
    return _raft(RAFT_SMALL, pretrained=pretrained, **kwargs)
