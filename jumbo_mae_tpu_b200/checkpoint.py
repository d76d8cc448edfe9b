"""Parameter checkpoints in the reference's wire format (Flax msgpack).

The reference saves ``msgpack_serialize(unreplicate(state.params))`` to ``<name>-last.msgpack`` /
``<name>-best.msgpack`` (src/main_pretrain.py:80-88, src/utils.py:55-63) and fine-tuning reads such a file back
with ``flax.serialization.msgpack_restore`` (src/utils.py:150-152).  Flax is not importable here, so this module
restates the published wire format of ``flax.serialization`` (flax 0.8.x, unchanged since 0.3):

  * the pytree is a msgpack map of str -> (map | leaf);
  * an ndarray leaf is ``ExtType(1, packb((shape, dtype.name, C-order bytes), use_bin_type=True))``;
  * numpy scalars are ``ExtType(3, <same triple>)``, Python complex ``ExtType(2, packb((re, im)))``;
  * arrays above 2**30 bytes are split into a map
    ``{"__msgpack_chunked_array__": True, "shape": [...], "chunks": {"0": ext, "1": ext, ...}}`` of flat pieces.

With it, weights move between the reference and this build in both directions: ``state.params`` has the Flax
names and shapes (params.ParamStore.tree).  SURVEY.md section 8(f) rank 2; host-side only, no kernel involved.
"""
from __future__ import annotations

from typing import Any, Dict

import msgpack
import numpy as np

EXT_NDARRAY, EXT_NATIVE_COMPLEX, EXT_NPSCALAR = 1, 2, 3
MAX_CHUNK_SIZE = 2 ** 30
_CHUNKED = "__msgpack_chunked_array__"


def _ndarray_to_bytes(arr: np.ndarray) -> bytes:
    if arr.dtype.hasobject or arr.dtype.isalignedstruct:
        raise ValueError("object and structured arrays cannot be serialised")
    return msgpack.packb((list(arr.shape), arr.dtype.name, arr.tobytes("C")), use_bin_type=True)


def _ndarray_from_bytes(data: bytes) -> np.ndarray:
    shape, dtype_name, buffer = msgpack.unpackb(data, raw=True)
    name = dtype_name.decode() if isinstance(dtype_name, bytes) else dtype_name
    if name == "bfloat16":
        raise ValueError("bfloat16 leaves are not supported by this reader (the reference stores f32 parameters)")
    return np.frombuffer(buffer, dtype=np.dtype(name)).reshape(shape).copy()


def _ext_pack(x: Any):
    if isinstance(x, np.ndarray):
        return msgpack.ExtType(EXT_NDARRAY, _ndarray_to_bytes(x))
    if isinstance(x, np.generic):
        return msgpack.ExtType(EXT_NPSCALAR, _ndarray_to_bytes(np.asarray(x)))
    if isinstance(x, complex):
        return msgpack.ExtType(EXT_NATIVE_COMPLEX, msgpack.packb((x.real, x.imag)))
    return x


def _ext_unpack(code: int, data: bytes):
    if code == EXT_NDARRAY:
        return _ndarray_from_bytes(data)
    if code == EXT_NATIVE_COMPLEX:
        re, im = msgpack.unpackb(data)
        return complex(re, im)
    if code == EXT_NPSCALAR:
        return _ndarray_from_bytes(data)[()]
    return msgpack.ExtType(code, data)


def _tuple_to_dict(tpl) -> Dict[str, Any]:
    return {str(i): v for i, v in enumerate(tpl)}


def _dict_to_tuple(d: Dict[str, Any]):
    return tuple(d[str(i)] for i in range(len(d)))


def _chunk(arr: np.ndarray) -> Dict[str, Any]:
    chunksize = max(1, int(MAX_CHUNK_SIZE / arr.dtype.itemsize))
    flat = arr.reshape(-1)
    chunks = [flat[i:i + chunksize] for i in range(0, flat.size, chunksize)]
    return {_CHUNKED: True, "shape": _tuple_to_dict(arr.shape), "chunks": _tuple_to_dict(chunks)}


def _prepare(tree: Any) -> Any:
    """torch tensors / array-likes -> C-contiguous numpy leaves, oversize leaves chunked."""
    if isinstance(tree, dict):
        return {str(k): _prepare(v) for k, v in tree.items()}
    if hasattr(tree, "detach"):                      # torch.Tensor (possibly a strided device view)
        tree = tree.detach().to("cpu").contiguous().numpy()
    if isinstance(tree, np.ndarray):
        arr = tree if tree.flags.c_contiguous else np.ascontiguousarray(tree)   # (ascontiguousarray makes 0-d 1-d)
        return _chunk(arr) if arr.size * arr.dtype.itemsize > MAX_CHUNK_SIZE else arr
    return tree


def _unchunk(tree: Any) -> Any:
    if isinstance(tree, dict):
        if tree.get(_CHUNKED, False):
            shape = _dict_to_tuple(tree["shape"])
            return np.concatenate(_dict_to_tuple(tree["chunks"])).reshape(shape)
        return {k: _unchunk(v) for k, v in tree.items()}
    return tree


def msgpack_serialize(pytree: Any) -> bytes:
    """flax.serialization.msgpack_serialize for dict pytrees of arrays (src/main_pretrain.py:83)."""
    return msgpack.packb(_prepare(pytree), default=_ext_pack, strict_types=True)


def msgpack_restore(encoded: bytes) -> Any:
    """flax.serialization.msgpack_restore (src/utils.py:151): nested dicts of numpy arrays."""
    return _unchunk(msgpack.unpackb(encoded, ext_hook=_ext_unpack, raw=False, strict_map_key=False))


# ----------------------------------------------------------------------------- train-state helpers
def save_params(state, path: str) -> None:
    """Writes ``state.params`` as the reference does (``<name>-last.msgpack``, src/utils.py:55-63)."""
    data = msgpack_serialize(state.params)
    with open(path, "wb") as f:
        f.write(data)


def load_params(state, path: str) -> None:
    """Loads a parameter file written by the reference (or by ``save_params``) into ``state``: every leaf of the
    model must be present with the Flax shape; the bf16 compute copy is refreshed."""
    with open(path, "rb") as f:
        tree = msgpack_restore(f.read())
    if set(tree.keys()) == {"params"}:               # a whole Flax variable dict rather than its "params" entry
        tree = tree["params"]
    import torch

    def to_torch(t):
        return {k: to_torch(v) for k, v in t.items()} if isinstance(t, dict) else torch.from_numpy(np.asarray(t))
    state.store.load_tree(to_torch(tree))


def load_pretrained_params(state, path: str) -> int:
    """Fine-tuning from a pre-training checkpoint (``--pretrained-ckpt``, src/utils.py:150-202, called at
    src/finetuning.py:215-217): the leaves of the checkpoint's ``model`` sub-tree that the classification tree also has
    (patch embedding, CLS tokens, shared Jumbo MLP, every layer, final norm) replace the fresh initialisation; the
    head keeps its initial values, the checkpoint's decoder leaves are ignored.  (The reference's active code returns
    the restored tree as it is, which only works for a checkpoint that already has the classification tree; the
    overlap it prints is what is loaded here.)  Returns the number of ``model`` entries taken over."""
    import torch
    with open(path, "rb") as f:
        new = msgpack_restore(f.read())
    if set(new.keys()) == {"params"}:
        new = new["params"]
    mine = state.store.tree()
    overlap = sorted(set(new["model"]).intersection(mine["model"]))

    def merge(dst, src, path_):
        out = {}
        for k, v in dst.items():
            if isinstance(v, dict):
                out[k] = merge(v, src[k], path_ + (k,)) if k in src else v
            elif k in src:
                t = torch.from_numpy(np.asarray(src[k]))
                if tuple(t.shape) != tuple(v.shape):
                    raise ValueError(f"{'/'.join(path_ + (k,))}: checkpoint shape {tuple(t.shape)} != {tuple(v.shape)}")
                out[k] = t
            else:
                out[k] = v
        return out
    merged = dict(mine)
    merged["model"] = {k: (merge(v, new["model"][k], ("model", k)) if isinstance(v, dict) and k in overlap
                           else (torch.from_numpy(np.asarray(new["model"][k])) if k in overlap else v))
                       for k, v in mine["model"].items()}
    state.store.load_tree(merged)
    return len(overlap)


def save_state(state, path: str) -> None:
    """Everything a run needs to resume (the reference saves parameters only and cannot): parameters, both Adam
    moments, the step / micro-step counters, an in-flight gradient accumulator and the RNG stream seeds, as
    one msgpack file in the same leaf encoding."""
    tree = {
        "params": state.params,
        "opt_state": {"mu": state.opt_state["mu"], "nu": state.opt_state["nu"]},
        "step": np.int64(state.step), "micro_step": np.int64(state.micro_step),
        "rngs": {"mixup": np.int64(state.mixup_rng), "dropout": np.int64(state.dropout_rng),
                 "noise": np.int64(state.noise_rng)},
    }
    if state.micro_step > 0:
        tree["grad_accum"] = state.store.tree(state.store.g)
    noise_gen = getattr(state, "_noise_gen", None)
    if noise_gen is not None:                      # position of this rank's mask-noise stream (Philox counter state)
        st = noise_gen.bit_generator.state
        tree["noise_stream"] = {"counter": np.asarray(st["state"]["counter"], dtype=np.uint64),
                                "key": np.asarray(st["state"]["key"], dtype=np.uint64),
                                "buffer": np.asarray(st["buffer"], dtype=np.uint64),
                                "buffer_pos": np.int64(st["buffer_pos"]), "has_uint32": np.int64(st["has_uint32"]),
                                "uinteger": np.int64(st["uinteger"])}
    with open(path, "wb") as f:
        f.write(msgpack_serialize(tree))


def load_state(state, path: str) -> None:
    """Inverse of ``save_state`` into a state built by ``create_train_state`` with the same arguments."""
    import torch
    with open(path, "rb") as f:
        tree = msgpack_restore(f.read())

    def to_torch(t):
        return {k: to_torch(v) for k, v in t.items()} if isinstance(t, dict) else torch.from_numpy(np.asarray(t))

    store = state.store
    store.load_tree(to_torch(tree["params"]))

    def fill(buf, src):
        def walk(dst, s):
            for k, v in dst.items():
                if isinstance(v, dict):
                    walk(v, s[k])
                else:
                    v.copy_(torch.from_numpy(np.asarray(s[k])).to(v.device).reshape(v.shape))
        walk(store.tree(buf), src)
    fill(store.m, tree["opt_state"]["mu"])
    fill(store.v, tree["opt_state"]["nu"])
    state.step, state.micro_step = int(tree["step"]), int(tree["micro_step"])
    if "grad_accum" in tree:
        fill(store.g, tree["grad_accum"])
    state.mixup_rng, state.dropout_rng = int(tree["rngs"]["mixup"]), int(tree["rngs"]["dropout"])
    state.noise_rng = int(tree["rngs"]["noise"])
    if "noise_stream" in tree:
        ns = tree["noise_stream"]
        gen = np.random.Generator(np.random.Philox(key=0))
        st = gen.bit_generator.state
        st["state"]["counter"] = np.asarray(ns["counter"], dtype=np.uint64)
        st["state"]["key"] = np.asarray(ns["key"], dtype=np.uint64)
        st["buffer"] = np.asarray(ns["buffer"], dtype=np.uint64)
        st["buffer_pos"], st["has_uint32"], st["uinteger"] = int(ns["buffer_pos"]), int(ns["has_uint32"]), int(ns["uinteger"])
        gen.bit_generator.state = st
        state._noise_gen = gen
