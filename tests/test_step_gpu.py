"""Whole-step parity on the GPU: forward activations, loss, every parameter gradient and
multi-step AdamW trajectories of the kernel schedule (jumbo_mae_tpu_b200.engine) against
the f32 CPU oracle on identical parameters, images and noise.

Tolerances (bf16 operands / f32 accumulation vs an f32 reference):
  masking indices                      exact
  activations (relative L2)            <= 1e-2 encoder, <= 3e-2 decoder output
  loss (relative)                      <= 5e-3
  per-leaf gradient (relative L2)      <= 3e-2, cosine >= 0.999
  k-step parameter delta (relative L2) <= 5e-2 on leaves whose update is not sign-noise
"""
import math

import numpy as np
import pytest
import torch

from oracle import jumbo_mae_oracle as O

pytestmark = pytest.mark.gpu


def _cfgs():
    return {
        "tiny": (O.tiny(), 4),
        "tiny_normpix": (O.tiny(norm_pix_loss=True), 3),
        "mid": (O.MAEConfig(layers=2, dim=256, heads=4, patch_size=16, image_size=64, dec_layers=2, dec_dim=128,
                            dec_heads=4, warmup_steps=10, training_steps=100, train_batch_size=256), 8),
        "vitb_width": (O.MAEConfig(layers=2, dec_layers=2, warmup_steps=10, training_steps=100), 3),
        "vith_patch14": (O.MAEConfig(layers=1, dim=160, heads=2, patch_size=14, image_size=56, dec_layers=1,
                                     dec_dim=64, dec_heads=2, warmup_steps=10, training_steps=100), 5),
        # BASELINE.json configs[0] exactly: ViT-B/16, 12 + 8 layers, 224 px, batch 8 (error growth over 20 blocks,
        # the T = 199 / head_dim 32 decoder at depth, the batched Jumbo weight gradient over K = 12 * B)
        "config1_vitb16": (O.MAEConfig(), 8),
        # one ViT-H/14 block at its real width: d = 1280, head_dim 80, P = 588 (row pitch padded to 592),
        # T = 67 / 259, decoder 512 x 16 heads
        # --posemb learnable (src/modeling.py:115-118): embed/wpe is a trained [grid, grid, dim] parameter
        "tiny_learnable_wpe": (O.tiny(posemb="learnable"), 4),
        # --layerscale / --dec-layerscale (src/modeling.py:159-166,185-201): scale1/2/3 per block, values ~0.5 +- 0.2
        "tiny_layerscale": (O.tiny(layerscale=True, dec_layerscale=True), 4),
        "mid_layerscale": (O.MAEConfig(layers=2, dim=256, heads=4, patch_size=16, image_size=64, dec_layers=2, dec_dim=128,
                                       dec_heads=4, warmup_steps=10, training_steps=100, train_batch_size=256,
                                       layerscale=True, dec_layerscale=True), 8),
        "vith_width": (O.MAEConfig(layers=2, dim=1280, heads=16, patch_size=14, image_size=224, dec_layers=1,
                                   warmup_steps=10, training_steps=100), 2),
        # two ViT-L/16 blocks at their real width (BASELINE.json configs[2]): d = 1024, 16 heads of 64, decoder 512 x 16
        "vitl_width": (O.MAEConfig(layers=2, dim=1024, heads=16, dec_layers=1, warmup_steps=10, training_steps=100), 2),
    }


def _product_cfg(ocfg):
    from jumbo_mae_tpu_b200.config import MAEConfig
    return MAEConfig(**{f: getattr(ocfg, f) for f in MAEConfig.__dataclass_fields__ if hasattr(ocfg, f)})


def _setup(name, fused_head=None):
    from jumbo_mae_tpu_b200.engine import MAEEngine
    from jumbo_mae_tpu_b200.params import ParamStore
    ocfg, B = _cfgs()[name]
    cfg = _product_cfg(ocfg)
    state = O.create_state(ocfg, seed=0, randomize_all=True)
    images, noise = O.synthetic_batch(ocfg, B, seed=1, noise_seed=1001)
    store = ParamStore(cfg)
    store.load_tree(state.params)
    eng = MAEEngine(cfg, store, B, fused_head=fused_head)
    return ocfg, cfg, state, images, noise, store, eng


def _rel(a, b):
    return ((a - b).norm() / (b.norm() + 1e-12)).item()


ALL = ["tiny", "tiny_normpix", "mid", "vitb_width", "vith_patch14", "config1_vitb16", "vith_width",
       "tiny_learnable_wpe", "tiny_layerscale", "mid_layerscale", "vitl_width"]


@pytest.mark.parametrize("fused_head", [True, False])
@pytest.mark.parametrize("name", ALL)
def test_forward_loss_and_activations(name, fused_head):
    """fused_head: dec_norm + pixel head on the masked rows only with the loss in the GEMM epilogue (the default);
    False: head on every row -> f32 prediction -> jm_patch_loss."""
    ocfg, cfg, state, images, noise, store, eng = _setup(name, fused_head)
    assert eng.fused_head == fused_head
    collect = {}
    want = O.pretrain_forward(state.params, images, noise, ocfg, collect)["loss"].item()
    loss = eng.forward(images.cuda(), torch.from_numpy(noise).cuda(), want_grad=False)
    torch.cuda.synchronize()
    B = images.shape[0]
    assert np.array_equal(eng.ids_restore.cpu().numpy(), collect["ids_restore"].numpy())       # bit exact
    assert np.array_equal(eng.mask_row.cpu().numpy(), collect["mask"][0].numpy())
    acts = {
        "enc_in": eng.xa[0], f"enc_layer_{ocfg.layers - 1}": eng.xa[ocfg.layers],
        "dec_in": eng.za[0], f"dec_layer_{ocfg.dec_layers - 1}": eng.za[ocfg.dec_layers],
    }
    for key, got in acts.items():
        w = collect[key]
        r = _rel(got.cpu().reshape(w.shape), w)
        assert r <= (1e-2 if key.startswith("enc") else 3e-2), f"{key}: rel l2 {r}"
    if not fused_head:
        pred = eng.pred.cpu().reshape(B, -1, ocfg.patch_dim)[:, 3:]
        assert _rel(pred, collect["image_output"]) <= 3e-2
    assert abs(loss.item() - want) <= 5e-3 * abs(want), (loss.item(), want)


@pytest.mark.parametrize("name,fused_head", [(n, True) for n in ALL] + [("tiny_normpix", False), ("vith_patch14", False)])
def test_gradients_match_oracle(name, fused_head):
    from jumbo_mae_tpu_b200.params import tree_leaves
    ocfg, cfg, state, images, noise, store, eng = _setup(name, fused_head)
    want_loss, want = O.loss_and_grads(state.params, images, noise, ocfg)
    eng.forward(images.cuda(), torch.from_numpy(noise).cuda())
    eng.backward()
    eng.layerscale_finalize()        # d scale1 / d scale2 from the accumulated gradients (no-op without LayerScale)
    torch.cuda.synchronize()
    got = dict(tree_leaves(store.tree(store.g)))
    bad = []
    ref = dict(O.tree_leaves(want))
    for path, w in ref.items():
        g = got[path].detach().cpu().reshape(w.shape)
        if path[-2:] == ("wk", "bias"):
            # softmax is invariant to a shift of every key, so d loss / d(key bias) is identically zero:
            # both sides hold rounding noise only.  Bound it against the sibling value-bias gradient.
            scale = ref[path[:-2] + ("wv", "bias")].norm().item()
            if not (g.norm().item() <= 3e-2 * scale and w.norm().item() <= 1e-3 * scale):
                bad.append(("/".join(path), g.norm().item(), scale))
            continue
        r = _rel(g, w)
        cos = (g.flatten() @ w.flatten() / (g.norm() * w.norm() + 1e-20)).item()
        if not (r <= 3e-2 and cos >= 0.999):
            bad.append(("/".join(path), r, cos))
    assert not bad, bad


def test_backward_accumulates_and_rezeroes():
    ocfg, cfg, state, images, noise, store, eng = _setup("tiny")
    im, nz = images.cuda(), torch.from_numpy(noise).cuda()
    eng.forward(im, nz)
    eng.backward()
    g1 = store.g.clone()
    eng.forward(im, nz)
    eng.backward(zero_grads=False)                  # gradient accumulation: sum (src/pretraining.py:153)
    assert torch.allclose(store.g, 2 * g1, rtol=1e-4, atol=1e-7)
    eng.forward(im, nz)
    eng.backward()
    assert torch.allclose(store.g, g1, rtol=1e-4, atol=1e-7)


@pytest.mark.parametrize("name", ["tiny", "mid", "config1_vitb16", "tiny_learnable_wpe"])
def test_k_step_training_matches_oracle(name):
    from jumbo_mae_tpu_b200.engine import lr_schedule
    from jumbo_mae_tpu_b200.params import tree_leaves
    ocfg, cfg, state, images, noise, store, eng = _setup(name)
    im = images.cuda()
    k = 3 if name == "config1_vitb16" else 5     # the full-depth oracle step costs seconds on the host
    p0 = dict(O.tree_leaves(state.params))
    losses_o, losses_g = [], []
    for step in range(k):
        nz = np.random.default_rng(100 + step).random(ocfg.num_patches, dtype=np.float32)
        state, met, _ = O.training_step(state, images, nz)
        losses_o.append(met["loss"])
        loss = eng.forward(im, torch.from_numpy(nz).cuda())
        eng.backward()
        lr = lr_schedule(step, cfg)
        assert abs(lr - met["learning_rate"]) < 1e-12
        eng.adamw_step(step, lr)
        losses_g.append(loss.item())
    torch.cuda.synchronize()
    for a, b in zip(losses_g, losses_o):
        assert abs(a - b) <= 5e-3 * abs(b)
    got = dict(tree_leaves(store.tree()))
    worst = 0.0
    for path, w in O.tree_leaves(state.params):
        if path[-2:] == ("wk", "bias"):
            continue        # zero true gradient: Adam turns rounding noise into +-lr steps on both sides
        delta_w = w - p0[path]
        delta_g = got[path].detach().cpu().reshape(w.shape) - p0[path]
        # Early Adam steps move every weight by ~lr * sign(g): leaves whose gradient is below bf16
        # resolution of its neighbours flip sign at random; compare in aggregate (relative L2).
        r = _rel(delta_g, delta_w)
        worst = max(worst, r)
        assert r <= 0.25, ("/".join(path), r)
    leaves = [(p, w) for p, w in O.tree_leaves(state.params) if p[-2:] != ("wk", "bias")]
    all_w = torch.cat([(w - p0[p]).flatten() for p, w in leaves])
    all_g = torch.cat([(got[p].detach().cpu().reshape(w.shape) - p0[p]).flatten() for p, w in leaves])
    assert _rel(all_g, all_w) <= 5e-2
    # bf16 compute copy tracks the f32 masters
    assert torch.equal(store.p16.float(), store.p.to(torch.bfloat16).float())


def test_engine_rejects_wrong_batch_shape():
    from jumbo_mae_tpu_b200._lib import JmError
    ocfg, cfg, state, images, noise, store, eng = _setup("tiny")
    with pytest.raises(JmError):
        eng.forward(images[:2].cuda(), torch.from_numpy(noise).cuda())


def _tiny_args(**kw):
    import argparse
    base = dict(layers=2, dim=64, heads=2, labels=0, patch_size=8, image_size=32, image_mask_ratio=0.75,
                dec_layers=2, dec_dim=64, dec_heads=2, norm_pix_loss=False, init_seed=0, mixup_seed=0,
                dropout_seed=0, noise_seed=0, optimizer="adamw", learning_rate=1.5e-4, weight_decay=0.05,
                adam_b1=0.9, adam_b2=0.95, adam_eps=1e-8, lr_decay=1.0, clip_grad=0.0, grad_accum=1,
                train_batch_size=256, warmup_steps=10, training_steps=100, posemb="sincos2d", dropout=0.0,
                droppath=0.0, dec_dropout=0.0, dec_droppath=0.0, layerscale=False, dec_layerscale=False)
    base.update(kw)
    return argparse.Namespace(**base)


@pytest.mark.parametrize("accum", [1, 2])
def test_training_step_api_matches_oracle(accum):
    """create_train_state / training_step / validation_step (reference names, src/pretraining.py:125-270),
    through the CUDA-graph path, with host uint8 batches, against the oracle with the same parameters."""
    from jumbo_mae_tpu_b200.pretraining import create_train_state, training_step, validation_step
    from jumbo_mae_tpu_b200.params import tree_leaves
    state = create_train_state(_tiny_args(grad_accum=accum)).replicate()
    ocfg = O.tiny()
    ostate = O.create_state(ocfg, seed=0, randomize_all=True)
    state.store.load_tree(ostate.params)
    images, noise = O.synthetic_batch(ocfg, 4 * accum, seed=5, noise_seed=6)
    val = validation_step(state, (images[:4],), noise=noise)
    want_val = O.pretrain_forward(ostate.params, images[:4], noise, ocfg)["loss"].item()
    assert abs(val["loss"].item() - want_val) <= 5e-3 * want_val and val["num_samples"].item() == 4
    # reference semantics of accumulation: grads of micro-batches are summed then divided (:135,:153)
    grads = None
    for i in range(accum):
        _, g = O.loss_and_grads(ostate.params, images[4 * i:4 * i + 4], noise, ocfg)
        grads = g if grads is None else O.tree_map(lambda a, b: a + b, grads, g)
    grads = O.tree_map(lambda a: a / accum, grads)
    want_state, want_metrics, _ = O.training_step(ostate, images[:4], noise, grads_override=grads)
    for i in range(accum):
        state, metrics = training_step(state, (images[4 * i:4 * i + 4].pin_memory(),), noise=noise)
        assert state.step == (1 if i == accum - 1 else 0)
    assert abs(metrics["learning_rate"] - want_metrics["learning_rate"]) < 1e-12
    assert torch.isfinite(metrics["loss"]).item()
    got = dict(tree_leaves(state.params))
    p0 = dict(O.tree_leaves(ostate.params))
    leaves = [(p, w) for p, w in O.tree_leaves(want_state.params) if p[-2:] != ("wk", "bias")]
    dw = torch.cat([(w - p0[p]).flatten() for p, w in leaves])
    dg = torch.cat([(got[p].detach().cpu().reshape(w.shape) - p0[p]).flatten() for p, w in leaves])
    assert _rel(dg, dw) <= 5e-2
    assert set(state.opt_state) >= {"count", "mu", "nu", "hyperparams"}


def test_cuda_graph_path_equals_eager():
    ocfg, cfg, state, images, noise, store, eng = _setup("mid")
    im, nz = images.cuda(), torch.from_numpy(noise).cuda()
    loss_e = eng.forward(im, nz).clone()
    eng.backward()
    g_e = store.g.clone()
    loss_g = eng.run_graph(im, nz, zero_grads=True).clone()
    torch.cuda.synchronize()
    # identical kernels; only the order of the f32 reductions in L2 (split-K of the Jumbo MLP output in the
    # forward pass, weight-gradient accumulation in the backward pass) may differ from run to run
    assert abs(loss_e.item() - loss_g.item()) <= 1e-5 * abs(loss_e.item())
    # (a last-bit change of a forward value can flip bf16 roundings downstream: compare in the l2 norm)
    assert (store.g - g_e).norm().item() <= 2e-3 * g_e.norm().item()
    eng.run_graph(im, nz, zero_grads=False)
    assert (store.g - 2 * g_e).norm().item() <= 4e-3 * g_e.norm().item()


def test_segmented_backward_graphs_equal_whole_backward():
    """The data-parallel path replays forward + backward as one graph per segment (gradient all-reduces start
    between them); on one rank it must produce the same gradients as the single-graph path."""
    from jumbo_mae_tpu_b200.parallel import GradReducer
    ocfg, cfg, state, images, noise, store, eng = _setup("mid")
    im, nz = images.cuda(), torch.from_numpy(noise).cuda()
    eng.run_graph(im, nz, zero_grads=True)
    g_ref = store.g.clone()
    red = GradReducer(store)
    loss = eng.run_graph_overlapped(im, nz, True, red)
    torch.cuda.synchronize()
    assert (store.g - g_ref).norm().item() <= 2e-3 * g_ref.norm().item()   # L2 reduction order, see above
    assert torch.isfinite(loss).all()
    # decoder segment + encoder segments (layers L-1 .. 1) + the Jumbo segment + the tail (rest of layer 0, embedding)
    assert eng.num_bwd_segments() == 3 + (cfg.layers - 1 + eng.ENC_LAYERS_PER_SEGMENT - 1) // eng.ENC_LAYERS_PER_SEGMENT


def test_prefetched_batches_give_the_same_step():
    """TrainState.prefetch (copy stream + staging buffers) feeds training_step the same data as a direct call."""
    from jumbo_mae_tpu_b200.pretraining import create_train_state, training_step
    ocfg = O.tiny()
    ostate = O.create_state(ocfg, seed=0, randomize_all=True)
    images, noise = O.synthetic_batch(ocfg, 4, seed=5, noise_seed=6)
    host = [images.pin_memory(), torch.flip(images, dims=[0]).contiguous().pin_memory(), images.pin_memory()]

    def run(prefetch):
        state = create_train_state(_tiny_args()).replicate()
        state.store.load_tree(ostate.params)
        losses = []
        nxt = state.prefetch((host[0],)) if prefetch else (host[0],)
        for i in range(len(host)):
            cur = nxt
            state, m = training_step(state, cur, noise=noise)
            if i + 1 < len(host):
                nxt = state.prefetch((host[i + 1],)) if prefetch else (host[i + 1],)
            losses.append(m["loss"].item())
        return losses, state.store.p.clone()

    want, p_want = run(False)
    got, p_got = run(True)
    for a, b in zip(want, got):
        assert abs(a - b) <= 1e-5 * abs(a)
    assert (p_got - p_want).norm().item() <= 1e-3 * p_want.norm().item() * 1e-2 + 1e-6


def test_training_step_with_gradient_clipping_matches_oracle():
    """--clip-grad > 0: chain(clip_by_global_norm, adamw) (src/pretraining.py:243-244), threshold low enough to bind."""
    import dataclasses
    from jumbo_mae_tpu_b200.pretraining import create_train_state, training_step
    from jumbo_mae_tpu_b200.params import tree_leaves
    ocfg = dataclasses.replace(O.tiny(), clip_grad=0.05)
    ostate = O.create_state(ocfg, seed=0, randomize_all=True)
    images, noise = O.synthetic_batch(ocfg, 4, seed=5, noise_seed=6)
    _, grads = O.loss_and_grads(ostate.params, images, noise, ocfg)
    gnorm = torch.sqrt(sum((g.double() ** 2).sum() for _, g in O.tree_leaves(grads))).item()
    assert gnorm > 2 * ocfg.clip_grad                       # the clip is active in this test
    want_state, _, _ = O.training_step(ostate, images, noise)
    unclipped_state, _, _ = O.training_step(dataclasses.replace(ostate, cfg=dataclasses.replace(ocfg, clip_grad=0.0)),
                                            images, noise)
    state = create_train_state(_tiny_args(clip_grad=0.05)).replicate()
    state.store.load_tree(ostate.params)
    state, _ = training_step(state, (images.pin_memory(),), noise=noise)
    assert abs(state._sumsq.sqrt().item() - gnorm) <= 3e-2 * gnorm
    # Adam's first step moves every parameter by ~lr * sign(g) whatever the gradient scale, so the parameters barely
    # see the clip; the first moment does: mu = (1 - b1) * clipped g
    got_mu = dict(tree_leaves(state.opt_state["mu"]))
    leaves = [(p, w) for p, w in O.tree_leaves(want_state.mu) if p[-2:] != ("wk", "bias")]
    want = torch.cat([w.flatten() for _, w in leaves])
    got = torch.cat([got_mu[p].detach().cpu().reshape(w.shape).flatten() for p, w in leaves])
    unclipped = torch.cat([dict(O.tree_leaves(unclipped_state.mu))[p].flatten() for p, _ in leaves])
    assert _rel(got, want) <= 3e-2
    assert _rel(unclipped, want) > 0.5                      # and the oracle without the clip is far away


def test_adamw_by_segment_ranges_equals_whole_buffer_update():
    """engine.adamw_ranges over the backward segments' ranges (what a data-parallel step runs as each segment's
    all-reduce finishes) is the same update as one adamw_step over the flat buffers."""
    ocfg, cfg, state, images, noise, store, eng = _setup("mid")
    eng.forward(images.cuda(), torch.from_numpy(noise).cuda())
    eng.backward()
    p0, m0, v0 = store.p.clone(), store.m.clone(), store.v.clone()
    eng.adamw_step(3, 1e-3, grad_scale=0.5)
    want = (store.p.clone(), store.m.clone(), store.v.clone(), store.p16.clone())
    store.p.copy_(p0), store.m.copy_(m0), store.v.copy_(v0)
    covered = 0
    for k in range(eng.num_bwd_segments()):
        ranges = eng.segment_grad_ranges(k)
        covered += sum(e - s for s, e in ranges)
        eng.adamw_ranges(ranges, 3, 1e-3, grad_scale=0.5)
    assert covered == store.total
    for a, b in zip(want, (store.p, store.m, store.v, store.p16)):
        assert torch.equal(a, b)


def test_layerwise_lr_decay_scales_every_leaf_by_its_label():
    """--lr-decay < 1 (src/pretraining.py:235-242): same gradients, same moments, and every leaf's parameter step is
    the plain AdamW step times lr_decay ** (layers - get_layer_index(leaf)).  The oracle's AdamW with the same
    gradients gives the same parameters (f32 update arithmetic: rel 1e-6 of the step)."""
    import dataclasses
    ocfg, cfg, state, images, noise, store, eng = _setup("mid")
    eng.forward(images.cuda(), torch.from_numpy(noise).cuda())
    eng.backward()
    p0, m0, v0 = store.p.clone(), store.m.clone(), store.v.clone()
    eng.adamw_step(3, 1e-3, grad_scale=0.5)
    p1, m1, v1 = store.p.clone(), store.m.clone(), store.v.clone()
    store.p.copy_(p0), store.m.copy_(m0), store.v.copy_(v0)
    eng.adamw_step(3, 1e-3, grad_scale=0.5, lr_decay=0.75)
    assert torch.equal(store.m, m1) and torch.equal(store.v, v1)
    scales = set()
    for name, leaf in store.leaves.items():
        sl = slice(leaf.offset, leaf.offset + leaf.size)
        scale = O.layerwise_scale(tuple(name.split("/")), dataclasses.replace(ocfg, lr_decay=0.75))
        scales.add(scale)
        plain, decayed = (p1[sl] - p0[sl]).double(), (store.p[sl] - p0[sl]).double()
        # the steps are differences of f32 parameters: compare with an absolute slack of a few ulp of the parameter
        slack = 4 * torch.finfo(torch.float32).eps * p0[sl].abs().max().item()
        assert (decayed - scale * plain).abs().max().item() <= slack + 1e-6 * plain.abs().max().item(), name
        assert torch.equal(store.p16[sl], store.p[sl].to(torch.bfloat16)), name
    assert len(scales) == ocfg.layers + 1 and min(scales) == 0.75 ** ocfg.layers and max(scales) == 1.0
    # against the oracle's optimizer with the engine's own gradients
    from jumbo_mae_tpu_b200.params import tree_leaves
    grads = {}
    for path, g in tree_leaves(store.tree(store.g)):
        node = grads
        for k in path[:-1]:
            node = node.setdefault(k, {})
        node[path[-1]] = g.detach().cpu().clone() * 0.5
    dcfg = dataclasses.replace(ocfg, lr_decay=0.75)
    want = {}
    for path, p in O.tree_leaves(state.params):
        g = grads
        for k in path:
            g = g[k]
        newp, _, _ = O.adamw_leaf(p, g.reshape(p.shape), torch.zeros_like(p), torch.zeros_like(p), 3,
                                  1e-3 * O.layerwise_scale(path, dcfg), dcfg, decay=(path[-1] == "kernel"))
        want[path] = newp
    got = dict(tree_leaves(store.tree(store.p)))
    for path, w in want.items():
        step = (w - dict(O.tree_leaves(state.params))[path]).abs().max().item()
        assert (got[path].detach().cpu().reshape(w.shape) - w).abs().max().item() <= 1e-3 * step + 1e-7, path


def test_training_step_accepts_lr_decay():
    """create_train_state(--lr-decay 0.75) builds and steps; the embedding moves less than with --lr-decay 1."""
    from jumbo_mae_tpu_b200.pretraining import create_train_state, training_step
    ocfg = O.tiny()
    ostate = O.create_state(ocfg, seed=0, randomize_all=True)
    images, noise = O.synthetic_batch(ocfg, 4, seed=5, noise_seed=6)
    moved = {}
    for decay in (1.0, 0.5):
        state = create_train_state(_tiny_args(lr_decay=decay)).replicate()
        state.store.load_tree(ostate.params)
        before = state.store.f("model/embed/wte/kernel").clone()
        top = state.store.f("decoder_proj/kernel").clone()
        state, m = training_step(state, (images.pin_memory(),), noise=noise)
        moved[decay] = ((state.store.f("model/embed/wte/kernel") - before).norm().item(),
                        (state.store.f("decoder_proj/kernel") - top).norm().item())
    L = ocfg.layers
    assert abs(moved[0.5][0] / moved[1.0][0] - 0.5 ** L) <= 5e-3 * 0.5 ** L      # f32 rounding of p at steps of 1e-6
    assert abs(moved[0.5][1] / moved[1.0][1] - 1.0) <= 1e-4


@pytest.mark.parametrize("name", ["tiny", "mid"])
def test_grad_ckpt_recomputes_the_same_step(name):
    """--grad-ckpt (nn.remat of every layer, src/modeling.py:232,280): one shared set of activation buffers, every
    block's forward recomputed in front of its backward -- the same kernels on the same data, so loss and gradients
    agree up to the order of the f32 reductions (split-K reduce-adds, vector atomics of the parameter gradients)."""
    import dataclasses
    from jumbo_mae_tpu_b200.engine import MAEEngine
    from jumbo_mae_tpu_b200.params import ParamStore
    ocfg, cfg, state, images, noise, store, eng = _setup(name)
    dev_images, dev_noise = images.cuda(), torch.from_numpy(noise).cuda()
    eng.forward(dev_images, dev_noise)
    eng.backward()
    torch.cuda.synchronize()
    want_loss, want = eng.loss.clone(), store.g.clone()
    cfg2 = dataclasses.replace(cfg, grad_ckpt=True)
    store2 = ParamStore(cfg2)
    store2.load_tree(state.params)
    eng2 = MAEEngine(cfg2, store2, images.shape[0])
    assert eng2.remat and eng2.enc[0].qkv.data_ptr() == eng2.enc[-1].qkv.data_ptr()     # one set of buffers
    assert eng2.activation_bytes() < eng.activation_bytes()
    eng2.forward(dev_images, dev_noise)
    eng2.backward()
    torch.cuda.synchronize()
    assert abs(eng2.loss.item() - want_loss.item()) <= 2e-5 * abs(want_loss.item())
    assert _rel(store2.g, want) <= 2e-3      # (an f32 sum in another order can flip a bf16 rounding downstream)


@pytest.mark.parametrize("lr_decay", [1.0, 0.75])
def test_lamb_step_matches_oracle(lr_decay):
    """--optimizer lamb (modified_lamb, src/utils.py:124-139): after one backward pass, engine.lamb_step against the
    oracle's lamb_leaf applied to every leaf of the Flax tree with the engine's own gradients -- trust ratio per Flax
    leaf (wq / wk / wv separately although they are stored fused), plain Adam on the non-kernel leaves, the layer-wise
    lr table behind it, moments updated once, bf16 copy refreshed."""
    import dataclasses
    from jumbo_mae_tpu_b200.params import tree_leaves
    ocfg, cfg, state, images, noise, store, eng = _setup("tiny")
    eng.forward(images.cuda(), torch.from_numpy(noise).cuda())
    eng.backward()
    torch.cuda.synchronize()
    store.m.normal_(0, 1e-3)
    store.v.uniform_(0, 1e-6)
    params0 = {path: p.detach().cpu().clone() for path, p in tree_leaves(store.tree(store.p))}
    grads = {path: g.detach().cpu().clone() for path, g in tree_leaves(store.tree(store.g))}
    mus = {path: t.detach().cpu().clone() for path, t in tree_leaves(store.tree(store.m))}
    nus = {path: t.detach().cpu().clone() for path, t in tree_leaves(store.tree(store.v))}
    count, lr, gscale = 3, 2e-3, 0.5
    eng.lamb_step(count, lr, grad_scale=gscale, lr_decay=lr_decay)
    torch.cuda.synchronize()
    dcfg = dataclasses.replace(ocfg, lr_decay=lr_decay)
    got_p = dict(tree_leaves(store.tree(store.p)))
    got_m = dict(tree_leaves(store.tree(store.m)))
    ratios = []
    for path, p0 in params0.items():
        decay = path[-1] == "kernel"
        want_p, want_m, _ = O.lamb_leaf(p0, grads[path] * gscale, mus[path], nus[path], count,
                                        lr * O.layerwise_scale(path, dcfg), dcfg, decay=decay)
        step = (want_p - p0).abs().max().item()
        assert (got_p[path].detach().cpu() - want_p).abs().max().item() <= 2e-4 * step + 1e-8, path
        assert torch.allclose(got_m[path].detach().cpu(), want_m, rtol=1e-5, atol=1e-9), path
        if decay:
            plain, _, _ = O.adamw_leaf(p0, grads[path] * gscale, mus[path], nus[path], count,
                                       lr * O.layerwise_scale(path, dcfg), dcfg, decay=True)
            ratios.append(step / max((plain - p0).abs().max().item(), 1e-30))
    assert max(ratios) / min(ratios) > 1.5            # the trust ratios differ from leaf to leaf (and from 1)
    assert torch.equal(store.p16, store.p.to(torch.bfloat16))


def test_training_step_accepts_lamb():
    """create_train_state(--optimizer lamb) builds and steps (src/pretraining.py:45-46)."""
    from jumbo_mae_tpu_b200.pretraining import create_train_state, training_step
    ocfg = O.tiny()
    images, noise = O.synthetic_batch(ocfg, 4, seed=5, noise_seed=6)
    state = create_train_state(_tiny_args(optimizer="lamb")).replicate()
    before = state.store.p.clone()
    state, metrics = training_step(state, (images,), noise=noise)
    torch.cuda.synchronize()
    assert state.step == 1 and math.isfinite(metrics["loss"].item())
    assert not torch.equal(state.store.p, before)


@pytest.mark.parametrize("preset", ["vit_b16", "vit_l16"])
def test_full_size_step_is_the_mean_of_its_halves(preset):
    """Size-independent property at the presets' full shapes (12 + 8 layers of ViT-B/16, 24 + 8 of ViT-L/16, 224 px),
    where the oracle is too slow: loss and gradient of one batch equal the mean over its two halves (same mask), with
    the gradients accumulated over the two half-sized micro-steps as training does."""
    from jumbo_mae_tpu_b200.config import MAEConfig
    from jumbo_mae_tpu_b200.engine import MAEEngine
    from jumbo_mae_tpu_b200.params import ParamStore
    from jumbo_mae_tpu_b200.pretraining import init_params_
    kw = dict(vit_b16=dict(), vit_l16=dict(layers=24, dim=1024, heads=16))[preset]
    cfg = MAEConfig(**kw)
    store = ParamStore(cfg)
    init_params_(store, 3)
    gen = torch.Generator().manual_seed(11)
    for name, leaf in store.leaves.items():            # biases / LayerNorm / CLS away from their 0 / 1 initial values
        if not leaf.decay:
            store.f(name).add_(torch.randn(store.f(name).shape, generator=gen).cuda() * 0.02)
    store.refresh_bf16()
    B = 16
    images = torch.randint(0, 256, (B, 3, 224, 224), generator=gen, dtype=torch.uint8).cuda()
    noise = torch.rand(cfg.num_patches, generator=gen).cuda()
    whole = MAEEngine(cfg, store, B)
    whole.forward(images, noise)
    whole.backward()
    torch.cuda.synchronize()
    loss_whole, g_whole = whole.loss.item(), store.g.clone()
    del whole
    half = MAEEngine(cfg, store, B // 2)
    losses = []
    for i in range(2):
        half.forward(images[i * B // 2:(i + 1) * B // 2], noise)
        half.backward(zero_grads=(i == 0))
        torch.cuda.synchronize()
        losses.append(half.loss.item())
    assert abs(loss_whole - sum(losses) / 2) <= 2e-5 * abs(loss_whole)
    assert math.isfinite(loss_whole) and g_whole.abs().max().item() > 0
    assert _rel(store.g * 0.5, g_whole) <= 3e-3          # bf16 activations: the batched Jumbo products round differently


@pytest.mark.parametrize("base", ["tiny", "tiny_learnable_wpe", "tiny_normpix"])
def test_per_image_masking_matches_oracle(base):
    """cfg.per_image_masking (opt-in extension of random_masking: noise [B, N], one permutation per image): index tables
    bit-exact per image, loss and every gradient against the oracle with the same per-image noise."""
    import dataclasses
    from jumbo_mae_tpu_b200.engine import MAEEngine
    from jumbo_mae_tpu_b200.params import ParamStore, tree_leaves
    ocfg, B = _cfgs()[base]
    cfg = dataclasses.replace(_product_cfg(ocfg), per_image_masking=True)
    state = O.create_state(ocfg, seed=0, randomize_all=True)
    images, _ = O.synthetic_batch(ocfg, B, seed=1, noise_seed=1001)
    noise = np.random.default_rng(5).random((B, ocfg.num_patches), dtype=np.float32)
    store = ParamStore(cfg)
    store.load_tree(state.params)
    eng = MAEEngine(cfg, store, B)
    collect = {}
    want_fwd = O.pretrain_forward(state.params, images, noise, ocfg, collect)["loss"].item()
    want_loss, want = O.loss_and_grads(state.params, images, noise, ocfg)
    eng.forward(images.cuda(), torch.from_numpy(noise).cuda())
    eng.backward()
    torch.cuda.synchronize()
    assert np.array_equal(eng.ids_restore.cpu().numpy(), collect["ids_restore"].numpy())
    assert np.array_equal(eng.mask_row.cpu().numpy(), collect["mask"].numpy())
    assert len({tuple(r) for r in eng.ids_restore.cpu().numpy().tolist()}) == B          # the images really differ
    assert abs(eng.loss.item() - want_fwd) <= 5e-3 * abs(want_fwd)
    got = dict(tree_leaves(store.tree(store.g)))
    bad = []
    ref = dict(O.tree_leaves(want))
    for path, w in ref.items():
        g = got[path].detach().cpu().reshape(w.shape)
        if path[-2:] == ("wk", "bias"):
            continue
        r = _rel(g, w)
        cos = (g.flatten() @ w.flatten() / (g.norm() * w.norm() + 1e-20)).item()
        if not (r <= 3e-2 and cos >= 0.999):
            bad.append(("/".join(path), r, cos))
    assert not bad, bad
    with pytest.raises(Exception):                                     # a shared [N] noise row is refused
        eng.forward(images.cuda(), torch.from_numpy(noise[0]).cuda())


def test_training_step_draws_per_image_noise():
    from jumbo_mae_tpu_b200.pretraining import create_train_state, training_step, validation_step
    ocfg = O.tiny()
    images, _ = O.synthetic_batch(ocfg, 4, seed=5, noise_seed=6)
    state = create_train_state(_tiny_args(per_image_masking=True)).replicate()
    state, metrics = training_step(state, (images,))
    torch.cuda.synchronize()
    eng = state.engine(4)
    assert tuple(eng.ids_restore.shape) == (4, ocfg.num_patches) and math.isfinite(metrics["loss"].item())
    assert math.isfinite(validation_step(state, (images,))["loss"].item())
