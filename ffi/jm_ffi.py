"""Registers the XLA FFI handlers of ffi/jm_ffi.cc with JAX and wraps the step's hot ops as jax functions.

Import-gated: this module needs ``jax`` (>= 0.4.31: ``jax.ffi``) and a built ``ffi/libjm_ffi.so``; neither exists in
this image (no jax / jaxlib wheels, no network), so here it only documents the binding a maintainer of the reference
adds.  With it, ``src/utils_mae.py:random_masking`` becomes ``jm_ffi.random_masking`` etc., and ``training_step`` keeps
``jax.value_and_grad`` through the ``jax.custom_vjp`` pairs below.
"""
from __future__ import annotations

import ctypes
import os

TARGETS = [
    "JmMaskSort", "JmIndexSequence", "JmPatchify", "JmExtractPatches", "JmPatchLoss", "JmPatchMseLoss", "JmLayerNormFwd",
    "JmLayerNormBwd", "JmDense", "JmMaskedTargets", "JmLayerNormFwdRows", "JmLayerNormBwdRows", "JmHeadLoss",
    "JmDenseResidual", "JmDenseBwd", "JmAttentionFwd", "JmAttentionBwd", "JmFillCls",
    "JmUnshuffleFwd", "JmUnshuffleBwd", "JmColsum", "JmLamb", "JmLayerScaleGrad", "JmAdamW", "JmClipByGlobalNorm", "JmCastBf16", "JmSoftmaxCe",
]


def register(lib_path: str | None = None):
    """jax.ffi.register_ffi_target for every handler symbol; returns the loaded library."""
    import jax
    here = os.path.dirname(os.path.abspath(__file__))
    lib = ctypes.CDLL(lib_path or os.path.join(here, "libjm_ffi.so"))
    for name in TARGETS:
        jax.ffi.register_ffi_target(name, jax.ffi.pycapsule(getattr(lib, name)), platform="CUDA")
    return lib


def random_masking(x, noise, keep_len: int):
    """src/utils_mae.py:88-96 with the noise supplied as data: (kept, mask, ids_restore)."""
    import jax
    import jax.numpy as jnp
    L = noise.shape[0]
    ids_shuffle, ids_restore, mask_row = jax.ffi.ffi_call(
        "JmMaskSort", (jax.ShapeDtypeStruct((L,), jnp.int32), jax.ShapeDtypeStruct((L,), jnp.int32),
                       jax.ShapeDtypeStruct((L,), jnp.float32)))(noise, keep=jnp.int32(keep_len))
    kept = jax.ffi.ffi_call("JmIndexSequence", jax.ShapeDtypeStruct((x.shape[0], keep_len, x.shape[2]), jnp.float32))(
        x, ids_shuffle[:keep_len])
    return kept, jnp.broadcast_to(mask_row[None, :], (x.shape[0], L)), ids_restore


def attention(qkv, batch: int, heads: int, q_scale: float):
    """softmax(q k^T) v with a custom VJP pairing JmAttentionFwd / JmAttentionBwd (src/modeling.py:136-137)."""
    import jax
    import jax.numpy as jnp
    rows, ld = qkv.shape
    T, d = rows // batch, ld // 3

    @jax.custom_vjp
    def f(qkv):
        return fwd(qkv)[0]

    def fwd(qkv):
        out, lse = jax.ffi.ffi_call("JmAttentionFwd", (jax.ShapeDtypeStruct((rows, d), jnp.bfloat16),
                                                      jax.ShapeDtypeStruct((batch, heads, T), jnp.float32)))(
            qkv, batch=jnp.int32(batch), heads=jnp.int32(heads))
        return out, (qkv, out, lse)

    def bwd(res, dout):
        qkv, out, lse = res
        dqkv, _ = jax.ffi.ffi_call("JmAttentionBwd", (jax.ShapeDtypeStruct(qkv.shape, jnp.bfloat16),
                                                     jax.ShapeDtypeStruct((ld,), jnp.float32)),
                                   input_output_aliases={4: 1})(
            qkv, out, dout, lse, jnp.zeros((ld,), jnp.float32), batch=jnp.int32(batch), heads=jnp.int32(heads),
            q_scale=jnp.float32(q_scale))
        return (dqkv,)

    f.defvjp(fwd, bwd)
    return f(qkv)


def adamw(p, g, m, v, n_decay: int, **hyper):
    """optax.adamw + apply_updates over flat buffers; p, m, v are donated and updated in place."""
    import jax
    import jax.numpy as jnp
    shapes = (jax.ShapeDtypeStruct(p.shape, jnp.float32),) * 3 + (jax.ShapeDtypeStruct(p.shape, jnp.bfloat16),)
    return jax.ffi.ffi_call("JmAdamW", shapes, input_output_aliases={0: 0, 2: 1, 3: 2})(
        p, g, m, v, n_decay=jnp.int64(n_decay), **{k: jnp.float32(x) for k, x in hyper.items()})
