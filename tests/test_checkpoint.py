"""Flax-msgpack parameter checkpoints (jumbo_mae_tpu_b200/checkpoint.py): the wire format is pinned by a
hand-derived byte vector, because flax itself cannot be imported here."""
import numpy as np
import pytest
import torch

from jumbo_mae_tpu_b200 import checkpoint as C


def test_known_answer_bytes():
    # {"a": float32[3] = [0, 1, 2]} in flax.serialization's format, derived by hand from the msgpack spec:
    # fixmap(1) | fixstr "a" | ext8 len=25 type=1 | fixarray(3): fixarray(1)[3], fixstr(7) "float32", bin8 len=12
    want = bytes.fromhex("81a161c71901" "939103" "a7" + b"float32".hex() + "c40c" "00000000" "0000803f" "00000040")
    got = C.msgpack_serialize({"a": np.arange(3, dtype=np.float32)})
    assert got == want
    back = C.msgpack_restore(want)
    assert back["a"].dtype == np.float32 and back["a"].tolist() == [0.0, 1.0, 2.0]


def test_round_trip_nested_tree_and_dtypes():
    rng = np.random.default_rng(0)
    tree = {"model": {"layer_0": {"kernel": rng.standard_normal((3, 4, 5)).astype(np.float32),
                                  "bias": rng.standard_normal(5).astype(np.float32)},
                      "count": np.int32(7), "scalar0d": np.array(2.5, dtype=np.float64)},
            "flags": {"u8": np.arange(6, dtype=np.uint8).reshape(2, 3)}, "z": 1 + 2j, "n": 3, "s": "text"}
    back = C.msgpack_restore(C.msgpack_serialize(tree))
    assert np.array_equal(back["model"]["layer_0"]["kernel"], tree["model"]["layer_0"]["kernel"])
    assert back["model"]["layer_0"]["bias"].dtype == np.float32
    assert back["model"]["count"] == 7 and back["model"]["count"].dtype == np.int32
    assert back["model"]["scalar0d"].shape == () and float(back["model"]["scalar0d"]) == 2.5
    assert np.array_equal(back["flags"]["u8"], tree["flags"]["u8"])
    assert back["z"] == 1 + 2j and back["n"] == 3 and back["s"] == "text"


def test_torch_views_are_serialised_in_c_order():
    t = torch.arange(24, dtype=torch.float32).view(4, 6)[:, :4].unflatten(1, (2, 2))   # strided view
    back = C.msgpack_restore(C.msgpack_serialize({"w": t}))
    assert back["w"].shape == (4, 2, 2) and np.array_equal(back["w"], t.contiguous().numpy())


def test_chunked_leaves(monkeypatch):
    monkeypatch.setattr(C, "MAX_CHUNK_SIZE", 64)           # 16 floats per chunk
    a = np.arange(100, dtype=np.float32).reshape(10, 10)
    data = C.msgpack_serialize({"big": a, "small": a[:1, :4].copy()})
    raw = C.msgpack.unpackb(data, ext_hook=C._ext_unpack, raw=False)
    assert raw["big"]["__msgpack_chunked_array__"] is True and len(raw["big"]["chunks"]) == 7
    assert raw["big"]["shape"] == {"0": 10, "1": 10}
    back = C.msgpack_restore(data)
    assert np.array_equal(back["big"], a) and np.array_equal(back["small"], a[:1, :4])


def test_oracle_parameter_tree_round_trips():
    from oracle import jumbo_mae_oracle as O
    params = O.init_params(O.tiny(), seed=0)
    back = C.msgpack_restore(C.msgpack_serialize(params))
    a, b = dict(O.tree_leaves(params)), dict(O.tree_leaves(O.tree_map(torch.from_numpy, back)))
    assert a.keys() == b.keys()
    for k in a:
        assert tuple(a[k].shape) == tuple(b[k].shape) and torch.equal(a[k].float(), b[k])


@pytest.mark.gpu
def test_train_state_save_and_load(tmp_path):
    import argparse
    import bench
    from jumbo_mae_tpu_b200.pretraining import create_train_state
    args = bench.preset_args("vit_b16", 8)
    args.layers, args.dim, args.heads, args.dec_layers, args.dec_dim, args.dec_heads = 2, 64, 2, 1, 64, 2
    state = create_train_state(args)
    path = str(tmp_path / "tiny-last.msgpack")
    C.save_params(state, path)
    want = state.store.p.clone()
    state.store.p.mul_(0.5)
    state.store.refresh_bf16()
    C.load_params(state, path)
    assert torch.equal(state.store.p, want)
    assert torch.equal(state.store.p16, want.to(torch.bfloat16))
    tree = C.msgpack_restore(open(path, "rb").read())
    assert tree["model"]["layer_0"]["attn"]["wq"]["kernel"].shape == (64, 2, 32)        # Flax DenseGeneral shapes
    assert tree["model"]["embed"]["wte"]["kernel"].shape == (16, 16, 3, 64)


@pytest.mark.gpu
def test_resume_reproduces_the_run(tmp_path):
    """save_state / load_state: a run resumed after 2 steps continues exactly like the uninterrupted one (same
    parameters, moments, step counter and mask-noise stream)."""
    import bench
    from jumbo_mae_tpu_b200.pretraining import create_train_state, training_step
    args = bench.preset_args("vit_b16", 8)
    args.layers, args.dim, args.heads, args.dec_layers, args.dec_dim, args.dec_heads = 2, 64, 2, 1, 64, 2
    g = torch.Generator().manual_seed(0)
    batches = [torch.randint(0, 256, (8, 3, 224, 224), generator=g, dtype=torch.uint8) for _ in range(4)]
    a = create_train_state(args).replicate()
    for b in batches[:2]:
        a, _ = training_step(a, (b,))
    path = str(tmp_path / "resume.msgpack")
    C.save_state(a, path)
    losses_a = []
    for b in batches[2:]:
        a, m = training_step(a, (b,))
        losses_a.append(m["loss"].item())
    r = create_train_state(args).replicate()
    C.load_state(r, path)
    assert r.step == 2
    losses_r = []
    for b in batches[2:]:
        r, m = training_step(r, (b,))
        losses_r.append(m["loss"].item())
    for x, y in zip(losses_a, losses_r):
        assert abs(x - y) <= 1e-4 * abs(x)            # same masks (noise stream restored), same weights
    assert (a.store.p - r.store.p).norm().item() <= 1e-4 * a.store.p.norm().item()
