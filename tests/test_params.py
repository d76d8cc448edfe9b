"""CPU tests of the host-side parameter layout (jumbo_mae_tpu_b200/params.py): the Flax-shaped
view tree has the reference's names and shapes, views alias the flat buffer, and the
weight-decay boundary separates exactly the leaves named ``kernel``."""
import torch

from oracle import jumbo_mae_oracle as O
from jumbo_mae_tpu_b200.config import MAEConfig, vit_b16, vit_h14
from jumbo_mae_tpu_b200.engine import lr_schedule
from jumbo_mae_tpu_b200.params import ParamStore, tree_leaves


def _pc(ocfg):
    return MAEConfig(**{f: getattr(ocfg, f) for f in MAEConfig.__dataclass_fields__ if hasattr(ocfg, f)})


def test_tree_matches_reference_names_and_shapes():
    for ocfg in (O.tiny(), O.MAEConfig(layers=1, dim=160, heads=2, patch_size=14, image_size=56, dec_layers=1,
                                       dec_dim=64, dec_heads=2)):
        want = O.init_params(ocfg, 0)
        store = ParamStore(_pc(ocfg), device="cpu")
        got = dict(tree_leaves(store.tree()))
        ref = dict(O.tree_leaves(want))
        assert set(got) == set(ref)
        for path, w in ref.items():
            assert tuple(got[path].shape) == tuple(w.shape), path


def test_views_alias_flat_buffer_and_decay_boundary():
    ocfg = O.tiny()
    store = ParamStore(_pc(ocfg), device="cpu")
    tree = store.tree()
    for i, (path, leaf) in enumerate(tree_leaves(tree)):
        leaf.fill_(float(i + 1))
    flagged = torch.zeros(store.total, dtype=torch.bool)
    for i, (path, leaf) in enumerate(tree_leaves(tree)):
        idx = (store.p == float(i + 1)).nonzero().flatten()
        assert idx.numel() == leaf.numel(), path
        is_kernel = path[-1] == "kernel"
        assert bool((idx < store.n_decay).all()) == is_kernel and bool((idx >= store.n_decay).all()) != is_kernel, path
        flagged[idx] = True
    assert store.n_decay % 4 == 0 and store.total % 4 == 0
    # everything not covered by a leaf is padding and stays zero
    assert torch.equal(store.p[~flagged], torch.zeros(int((~flagged).sum())))


def test_param_count_vit_b():
    store = ParamStore(vit_b16(), device="meta")
    n = sum(l.rows * l.cols for l in store.leaves.values())
    assert n == O.count_params(O.init_params(O.tiny())) or n == 154_191_360
    h = ParamStore(vit_h14(), device="meta")
    assert h.ld("decoder_image_output/kernel") == 592       # 588 padded to a multiple of 8


def test_lr_schedule_matches_oracle():
    ocfg = O.vit_b16()
    cfg = _pc(ocfg)
    for count in (0, 1, 100, cfg.warmup_steps - 1, cfg.warmup_steps, cfg.warmup_steps + 1, 250000, cfg.training_steps,
                  cfg.training_steps + 5):
        assert lr_schedule(count, cfg) == O.lr_schedule(count, ocfg)


def test_backward_segment_ranges_partition_the_gradient_buffer():
    """Data-parallel overlap: the per-segment all-reduce ranges are disjoint and cover every leaf once."""
    from jumbo_mae_tpu_b200.engine import MAEEngine
    for cfg in (vit_b16(), _pc(O.tiny()), _pc(O.tiny(layers=1)), _pc(O.vit_l16())):
        store = ParamStore(cfg, device="meta")
        eng = MAEEngine.__new__(MAEEngine)          # no device buffers: only the host-side planning is exercised
        eng.cfg, eng.ps = cfg, store
        seen = []
        for k in range(eng.num_bwd_segments()):
            seen += eng.segment_grad_ranges(k)
        seen.sort()
        assert all(a[1] <= b[0] for a, b in zip(seen, seen[1:]))
        assert sum(e - s for s, e in seen) == store.total
        last = eng.num_bwd_segments() - 1
        layers = [i for k in range(1, last) for i in eng._enc_segment_layers(k)]
        assert layers == list(reversed(range(cfg.layers)))
        # the shared Jumbo MLP is a segment of its own, layer 0 and the patch embedding go with the tail
        names = lambda k: {n for n, l in store.leaves.items() if any(s <= l.offset < e for s, e in eng.segment_grad_ranges(k))}
        assert names(last - 1) and all(n.startswith("model/jumbo_mlp/") for n in names(last - 1))
        assert all(n.startswith(("model/layer_0/", "model/embed/", "model/cls_tokens")) for n in names(last))


def test_classification_tree_and_segment_ranges():
    """labels > 0: no decoder leaves, LinearCLS head under model/head/Dense_0 (src/modeling.py:209-219,236-240); the
    classification engine's backward segments still partition the flat gradient buffer."""
    import dataclasses
    from jumbo_mae_tpu_b200.finetune_engine import FinetuneEngine
    cfg = dataclasses.replace(vit_b16(), labels=1000)
    assert cfg.keep == cfg.num_patches and cfg.num_masked == 0
    store = ParamStore(cfg, device="meta")
    names = set(store.leaves)
    assert {"model/head/Dense_0/kernel", "model/head/Dense_0/bias"} <= names
    assert not any(n.startswith(("decoder", "image_mask")) for n in names)
    head = store.leaves["model/head/Dense_0/kernel"]
    assert (head.rows, head.cols) == (3 * 768, 1000) and head.decay
    eng = FinetuneEngine.__new__(FinetuneEngine)
    eng.cfg, eng.ps = cfg, store
    seen = sorted(r for k in range(eng.num_bwd_segments()) for r in eng.segment_grad_ranges(k))
    assert all(a[1] <= b[0] for a, b in zip(seen, seen[1:]))
    assert sum(e - s for s, e in seen) == store.total


def test_layerscale_leaves_follow_the_reference_tree():
    """--layerscale / --dec-layerscale: scale1/2/3 vectors beside the norms (src/modeling.py:159-162,185-189), no weight decay
    (the mask keeps leaves named ``kernel``), layer-wise lr label of their layer."""
    from jumbo_mae_tpu_b200.params import get_layer_index
    ocfg = O.tiny(layerscale=True, dec_layerscale=True)
    store = ParamStore(_pc(ocfg), device="cpu")
    got = dict(tree_leaves(store.tree()))
    ref = dict(O.tree_leaves(O.init_params(ocfg, 0)))
    assert set(got) == set(ref)
    J, d, dd = ocfg.num_cls_tokens, ocfg.dim, ocfg.dec_dim
    assert tuple(got[("model", "layer_1", "scale3")].shape) == (J * d,)
    assert tuple(got[("decoder_model", "dec_layer_0", "scale2")].shape) == (dd,)
    assert ("decoder_model", "dec_layer_0", "scale3") not in got
    for name, leaf in store.leaves.items():
        if name.rsplit("/", 1)[-1].startswith("scale") and name.rsplit("/", 1)[-1] != "scale":
            assert not leaf.decay and leaf.offset >= store.n_decay
    assert get_layer_index("model/layer_1/scale2", ocfg.layers) == 2


def test_lamb_tables_cover_every_kernel_leaf_once_with_wqkv_split():
    """params.lamb_tables: one group per ``kernel`` leaf of the FLAX tree (the fused wqkv matrix gives three), chunks
    consecutive per group and covering every row exactly once."""
    ocfg = O.tiny()
    store = ParamStore(_pc(ocfg), device="cpu")
    groups, chunks = store.lamb_tables(chunk_elems=1000)
    flax_kernels = [p for p, _ in O.tree_leaves(O.init_params(ocfg, 0)) if p[-1] == "kernel"]
    assert groups.shape == (len(flax_kernels), 7) and chunks.shape[1] == 3
    covered = torch.zeros(store.total, dtype=torch.int32)
    for gi, (off, rows, ld, col0, ncols, c0, nc) in enumerate(groups.tolist()):
        mine = chunks[c0:c0 + nc]
        assert bool((mine[:, 0] == gi).all())
        assert mine[:, 2].sum().item() == rows and mine[0, 1].item() == 0
        assert torch.equal(mine[1:, 1], (mine[:-1, 1] + mine[:-1, 2]))
        for r in range(rows):
            covered[off + r * ld + col0: off + r * ld + col0 + ncols] += 1
    kernel_elems = sum(l.rows * l.cols for l in store.leaves.values() if l.decay)
    assert int((covered == 1).sum()) == kernel_elems and int(covered.max()) == 1
    assert bool((covered[store.n_decay:] == 0).all())
    # restricted to one layer-wise lr label
    g1, _ = store.lamb_tables(lambda n: n.startswith("model/layer_0/"))
    assert g1.shape[0] == 6            # wq, wk, wv, wo, w1, w2
    assert store.lamb_tables(lambda n: False) == (None, None)
