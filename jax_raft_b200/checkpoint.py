"""Checkpoint I/O in the reference's formats.

* ``convert_checkpoint(torch_checkpoint, output_file)`` mirrors
  ``/root/reference/scripts/convert_checkpoint.py:35-56``: a torchvision RAFT ``state_dict`` (``.pth``)
  becomes the Flax ``{params, batch_stats}`` tree, written as msgpack.
* ``load_msgpack(path)`` reads such a file — including the ones the reference downloads
  (``model.py:17-21``, ``684-689``: ``raft_large_C_T_SKHT_V2-ff5fadd5.msgpack`` ...) — into the
  ``variables`` tree ``model.apply`` takes.

The byte format is flax.serialization's: a msgpack map whose ndarray leaves are ``ExtType(1,
packb((shape, dtype_name, raw_bytes)))``.  flax itself is not installed on this image, so the format
is implemented from its published definition and round-trip tested here; reading a file written by
real flax could not be tested offline.
"""

from __future__ import annotations

from typing import Any, Dict

import msgpack
import numpy as np

from .model import torch_state_dict_from_variables, variables_from_torch_state_dict

_EXT_NDARRAY = 1
_EXT_NATIVE_COMPLEX = 2
_EXT_NPSCALAR = 3


def _pack_ext(x):
    if isinstance(x, np.ndarray):
        return msgpack.ExtType(_EXT_NDARRAY, msgpack.packb((x.shape, x.dtype.name, x.tobytes("C")), use_bin_type=True))
    if isinstance(x, np.generic):
        a = np.asarray(x)
        return msgpack.ExtType(_EXT_NPSCALAR, msgpack.packb((a.shape, a.dtype.name, a.tobytes("C")), use_bin_type=True))
    raise TypeError(f"cannot serialise {type(x)}")


def _unpack_ext(code, data):
    if code in (_EXT_NDARRAY, _EXT_NPSCALAR):
        shape, dtype_name, buf = msgpack.unpackb(data, raw=True)
        if isinstance(dtype_name, bytes):
            dtype_name = dtype_name.decode()
        arr = np.frombuffer(buf, dtype=np.dtype(dtype_name)).reshape(tuple(shape), order="C")
        return arr.copy() if code == _EXT_NDARRAY else arr[()]
    if code == _EXT_NATIVE_COMPLEX:
        re, im = msgpack.unpackb(data)
        return complex(re, im)
    return msgpack.ExtType(code, data)


def to_bytes(variables: Dict[str, Any]) -> bytes:
    """flax.serialization.to_bytes for a tree of numpy arrays."""

    def norm(t):
        if isinstance(t, dict):
            return {str(k): norm(v) for k, v in t.items()}
        return np.ascontiguousarray(np.asarray(t))

    return msgpack.packb(norm(variables), default=_pack_ext, strict_types=True, use_bin_type=True)


def from_bytes(data: bytes) -> Dict[str, Any]:
    """flax.serialization.msgpack_restore: bytes -> nested dict of numpy arrays."""
    return msgpack.unpackb(data, ext_hook=_unpack_ext, raw=False, strict_map_key=False)


def save_msgpack(variables: Dict[str, Any], path: str) -> None:
    with open(path, "wb") as f:
        f.write(to_bytes(variables))


def load_msgpack(path: str) -> Dict[str, Any]:
    with open(path, "rb") as f:
        tree = from_bytes(f.read())
    if "params" not in tree:
        raise ValueError(f"{path}: not a RAFT checkpoint (no 'params' collection)")
    tree.setdefault("batch_stats", {})
    return tree


def convert_checkpoint(torch_checkpoint: str, output_file: str) -> Dict[str, Any]:
    """scripts/convert_checkpoint.py:35-56 — torch ``state_dict`` file -> msgpack of the Flax tree."""
    import torch

    assert output_file.endswith(".msgpack")
    state_dict = torch.load(torch_checkpoint, map_location="cpu")
    variables = variables_from_torch_state_dict(state_dict)
    save_msgpack(variables, output_file)
    return variables


def export_torch_checkpoint(variables: Dict[str, Any], output_file: str) -> None:
    """Inverse direction: the Flax tree as a torchvision-loadable ``state_dict`` file."""
    import torch

    torch.save(torch_state_dict_from_variables(variables), output_file)


if __name__ == "__main__":  # same CLI as the reference script (convert_checkpoint.py:59-68)
    import sys

    if len(sys.argv) != 3:
        print("Usage: python -m jax_raft_b200.checkpoint <input_file.pth> <output_file.msgpack>")
        sys.exit(1)
    convert_checkpoint(sys.argv[1], sys.argv[2])
