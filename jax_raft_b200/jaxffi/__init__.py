"""JAX side of the drop-in: ``jax.ffi`` registration of the raftk handlers and a ``CorrBlock`` /
``upsample_flow`` pair for ``jax_raft.raft_large(corr_block=...)``.

STATUS: source-complete, never executed on the build image — ``jax``, ``jaxlib`` and ``flax`` are not
installed there and cannot be.  Importing this module without jax raises
ImportError with that explanation; nothing else in the package depends on it.

Usage where jax is installed::

    from jax_raft_b200 import jaxffi
    jaxffi.build()                      # g++ raftk_ffi.cc against jax.ffi.include_dir(), links libraftk.so
    jaxffi.register()
    model, variables = jax_raft.raft_large(pretrained=True, corr_block=jaxffi.CorrBlock(4, 4))
    jaxffi.patch_upsample_flow()        # jax_raft.model.upsample_flow has no injection keyword (model.py:69, 508)
    flow = model.apply(variables, image1, image2, train=False, num_flow_updates=32)[-1]
"""

from __future__ import annotations

import ctypes
import os
import subprocess

try:
    import jax
    import jax.numpy as jnp
    import numpy as np
except ImportError as exc:  # pragma: no cover - jax is absent on the build image
    raise ImportError(
        "jax_raft_b200.jaxffi needs jax/jaxlib (absent on the build image: no wheels, no network). "
        "The torch-hosted API in jax_raft_b200 works without it."
    ) from exc

_HERE = os.path.dirname(os.path.abspath(__file__))
_PKG = os.path.dirname(_HERE)
_ROOT = os.path.dirname(_PKG)
FFI_LIB = os.path.join(_PKG, "lib", "libraftk_ffi.so")

_TARGETS = {
    "raftk_corr_pyramid": "RaftkCorrPyramid",
    "raftk_corr_lookup": "RaftkCorrLookup",
    "raftk_convex_upsample": "RaftkConvexUpsample",
    "raftk_grid_sample": "RaftkGridSample",
    "raftk_altcorr_pool": "RaftkAltcorrPool",
    "raftk_altcorr_lookup": "RaftkAltcorrLookup",
}


def build(cuda_home: str = "/usr/local/cuda") -> str:
    """Compile raftk_ffi.cc against the headers jaxlib ships and link it to libraftk.so."""
    from .. import build as _build

    _build.build()
    cmd = ["g++", "-O2", "-std=c++17", "-fPIC", "-shared", f"-I{jax.ffi.include_dir()}", f"-I{_ROOT}/include",
           f"-I{cuda_home}/include", os.path.join(_HERE, "raftk_ffi.cc"), "-o", FFI_LIB,
           f"-L{_PKG}/lib", "-lraftk", f"-Wl,-rpath,{_PKG}/lib"]
    res = subprocess.run(cmd, capture_output=True, text=True)
    if res.returncode != 0:
        raise RuntimeError(f"g++ failed:\n{res.stderr}")
    return FFI_LIB


def register() -> None:
    lib = ctypes.CDLL(FFI_LIB)
    for target, symbol in _TARGETS.items():
        jax.ffi.register_ffi_target(target, jax.ffi.pycapsule(getattr(lib, symbol)), platform="CUDA")


def _scratch_bytes(b, nq, h, w, d) -> int:
    lib = ctypes.CDLL(os.path.join(_PKG, "lib", "libraftk.so"))
    lib.raftk_corr_pyramid_scratch_bytes.restype = ctypes.c_size_t
    return int(lib.raftk_corr_pyramid_scratch_bytes(b, nq, h, w, d))


class CorrBlock:
    """Duck-typed like jax_raft.model.CorrBlock (model.py:403-481): ``out_channels``,
    ``build_pyramid``, ``index_pyramid``; plain Python class, not an nn.Module."""

    def __init__(self, num_levels: int = 4, radius: int = 4):
        assert num_levels == 4, "the FFI handlers are bound for the reference's 4 levels"
        self.num_levels = num_levels
        self.radius = radius
        self.out_channels = num_levels * (2 * radius + 1) ** 2

    def build_pyramid(self, fmap1, fmap2):
        assert fmap1.shape == fmap2.shape, "Input feature maps should have the same shapes"
        b, h, w, d = fmap2.shape
        min_fmap_size = 2 * (2 ** (self.num_levels - 1))
        assert not any(s < min_fmap_size for s in (h, w)), "Feature maps are too small to be down-sampled by the correlation pyramid."
        q = b * h * w
        outs = [jax.ShapeDtypeStruct((q, h >> l, w >> l, 1), jnp.float32) for l in range(4)]
        outs.append(jax.ShapeDtypeStruct((max(_scratch_bytes(b, h * w, h, w, d), 1),), jnp.uint8))
        *levels, _scratch = jax.ffi.ffi_call("raftk_corr_pyramid", outs)(fmap1, fmap2)
        return levels

    def index_pyramid(self, corr_pyramid, centroids_coords):
        assert centroids_coords.shape[-1] == 2
        b, h, w, _ = centroids_coords.shape
        out = jax.ShapeDtypeStruct((b, h, w, self.out_channels), jnp.float32)
        return jax.ffi.ffi_call("raftk_corr_lookup", out)(*corr_pyramid, centroids_coords, radius=np.int32(self.radius))


class AltCorrBlock(CorrBlock):
    """On-the-fly variant: the opaque pyramid is (fmap1, fmap2 pooled 0..3 times)."""

    def build_pyramid(self, fmap1, fmap2):
        assert fmap1.shape == fmap2.shape, "Input feature maps should have the same shapes"
        b, h, w, d = fmap2.shape
        outs = [jax.ShapeDtypeStruct((b, h >> l, w >> l, d), jnp.float32) for l in range(1, 4)]
        pooled = jax.ffi.ffi_call("raftk_altcorr_pool", outs)(fmap2)
        return (fmap1, [fmap2, *pooled])

    def index_pyramid(self, corr_pyramid, centroids_coords):
        fmap1, levels = corr_pyramid
        b, h, w, _ = centroids_coords.shape
        out = jax.ShapeDtypeStruct((b, h, w, self.out_channels), jnp.float32)
        return jax.ffi.ffi_call("raftk_altcorr_lookup", out)(fmap1, *levels, centroids_coords, radius=np.int32(self.radius))


def upsample_flow(flow, up_mask=None, factor: int = 8):
    """model.py:69-98 with the mask branch on the sm_100a kernel; the bilinear branch defers to the
    reference."""
    if up_mask is None or factor != 8 or flow.shape[-1] != 2:
        import jax_raft.model as ref

        return ref._reference_upsample_flow(flow, up_mask, factor)
    b, h, w, _ = flow.shape
    assert up_mask.shape == (b, h, w, 9 * factor * factor)
    out = jax.ShapeDtypeStruct((b, 8 * h, 8 * w, 2), jnp.float32)
    return jax.ffi.ffi_call("raftk_convex_upsample", out)(flow, up_mask)


def patch_upsample_flow() -> None:
    """Swap jax_raft.model.upsample_flow (called at model.py:508) for the kernel-backed one."""
    import jax_raft.model as ref

    if not hasattr(ref, "_reference_upsample_flow"):
        ref._reference_upsample_flow = ref.upsample_flow
    ref.upsample_flow = upsample_flow


def grid_sample(img, grid):
    """model.py:24-34."""
    n, gh, gw, _ = grid.shape
    out = jax.ShapeDtypeStruct((n, gh, gw, img.shape[-1]), jnp.float32)
    return jax.ffi.ffi_call("raftk_grid_sample", out)(img, grid)
