"""CPU tests that pin the oracle (oracle/jumbo_mae_oracle.py).

The reference ships no tests, so these are hand-derivable
known-answer cases, one named test per assumed third-party default, plus a check
against the committed fixture tests/golden/tiny_step.npz.
"""
import math
import os

import numpy as np
import pytest
import torch

from oracle import jumbo_mae_oracle as O

GOLDEN = os.path.join(os.path.dirname(__file__), "golden", "tiny_step.npz")


def test_flops_and_param_counts_match_survey():
    fwd, step = O.flops_per_image(O.vit_b16())
    assert abs(fwd / 1e9 - 20.491) < 1e-3 and abs(step / 1e9 - 61.417) < 1e-3
    fwd, step = O.flops_per_image(O.vit_l16())
    assert abs(fwd / 1e9 - 45.001) < 1e-3 and abs(step / 1e9 - 134.925) < 1e-3
    fwd, step = O.flops_per_image(O.vit_h14())
    assert abs(fwd / 1e9 - 104.509) < 1e-3 and abs(step / 1e9 - 313.430) < 1e-3


def test_masking_known_answer():
    # src/utils_mae.py:90-96 with a hand-checkable noise vector
    noise = np.array([0.9, 0.1, 0.5, 0.3, 0.7, 0.2, 0.8, 0.4], dtype=np.float32)
    ids_shuffle, ids_restore, mask = O.masking_from_noise(noise, 2)
    assert ids_shuffle.tolist() == [1, 5, 3, 7, 2, 4, 6, 0]
    assert ids_restore.tolist() == [7, 0, 4, 2, 5, 1, 6, 3]
    assert mask.tolist() == [1, 0, 1, 1, 1, 0, 1, 1]
    assert ids_shuffle.dtype == np.int32 and ids_restore.dtype == np.int32


def test_masking_is_stable_on_ties():
    noise = np.array([0.5, 0.25, 0.5, 0.25, 0.5], dtype=np.float32)
    ids_shuffle, ids_restore, _ = O.masking_from_noise(noise, 2)
    assert ids_shuffle.tolist() == [1, 3, 0, 2, 4]          # equal keys keep index order
    assert ids_restore[ids_shuffle].tolist() == [0, 1, 2, 3, 4]


def test_random_masking_shapes_and_gather():
    x = torch.arange(2 * 8 * 3, dtype=torch.float32).reshape(2, 8, 3)
    noise = np.array([0.9, 0.1, 0.5, 0.3, 0.7, 0.2, 0.8, 0.4], dtype=np.float32)
    kept, mask, ids_restore = O.random_masking(x, noise, 2)
    assert kept.shape == (2, 2, 3) and mask.shape == (2, 8) and ids_restore.shape == (8,)
    assert torch.equal(kept, x[:, [1, 5]])
    assert torch.equal(mask[0], mask[1])                    # ONE permutation per call
    assert mask.sum().item() == 2 * 6


def test_extract_merge_patches_roundtrip():
    x = torch.randn(2, 32, 32, 3)
    p = O.extract_patches(x, 8)
    assert p.shape == (2, 16, 192)
    assert torch.equal(O.merge_patches(p, 8), x)
    # (ph, pw, c) order: element [0, patch(1,2), (3,4,c=1)] is pixel (8+3, 16+4, 1)
    assert p[0, 1 * 4 + 2, (3 * 8 + 4) * 3 + 1] == x[0, 11, 20, 1]


def test_patch_mse_all_ones_mask_is_plain_mse():
    o, t = torch.randn(3, 16, 12), torch.randn(3, 16, 12)
    assert torch.allclose(O.patch_mse_loss(o, t), ((o - t) ** 2).mean(), atol=1e-6)
    mask = torch.zeros(3, 16)
    mask[:, ::4] = 1
    want = ((o - t) ** 2).mean(-1)[:, ::4].mean()
    assert torch.allclose(O.patch_mse_loss(o, t, mask), want, atol=1e-6)


def test_sincos_known_values():
    e = O.fixed_sincos2d_embeddings(3, 4, 8)
    assert e.shape == (3, 4, 8)
    f = [1.0, 1.0 / 10000.0]                                # linspace(0,1,2) includes the endpoint
    a, b = 2, 3                                             # first axis -> last two groups
    want = [math.sin(b * f[0]), math.sin(b * f[1]), math.cos(b * f[0]), math.cos(b * f[1]),
            math.sin(a * f[0]), math.sin(a * f[1]), math.cos(a * f[0]), math.cos(a * f[1])]
    assert np.allclose(e[a, b].numpy(), want, atol=1e-6)


def test_layernorm_default_eps_and_biased_variance():
    x = torch.tensor([[1.0, 2.0, 3.0, 6.0]])
    y = O.layer_norm(x, torch.ones(4), torch.zeros(4))
    var = ((x - 3.0) ** 2).mean()                           # biased
    assert torch.allclose(y, (x - 3.0) / torch.sqrt(var + 1e-6), atol=1e-6)
    assert torch.allclose(y, torch.nn.functional.layer_norm(x, (4,), eps=1e-6), atol=1e-6)


def test_gelu_is_tanh_approximation():
    x = torch.linspace(-4, 4, 33)
    assert torch.allclose(O.gelu_tanh(x), torch.nn.functional.gelu(x, approximate="tanh"), atol=1e-6)
    assert not torch.allclose(O.gelu_tanh(x), torch.nn.functional.gelu(x), atol=1e-5)


def test_attention_q_scale_after_bias():
    torch.manual_seed(0)
    d, h = 8, 2
    p = {n: {"kernel": torch.randn(d, h, d // h), "bias": torch.randn(h, d // h)} for n in ("wq", "wk", "wv")}
    p["wo"] = {"kernel": torch.randn(h, d // h, d), "bias": torch.randn(d)}
    x = torch.randn(1, 5, d)
    y = O.attention(x, p, h)
    q = (x @ p["wq"]["kernel"].reshape(d, d) + p["wq"]["bias"].reshape(d)) / math.sqrt(d // h)
    k = x @ p["wk"]["kernel"].reshape(d, d) + p["wk"]["bias"].reshape(d)
    v = x @ p["wv"]["kernel"].reshape(d, d) + p["wv"]["bias"].reshape(d)
    outs = []
    for i in range(h):
        s = slice(i * d // h, (i + 1) * d // h)
        outs.append(torch.softmax(q[0, :, s] @ k[0, :, s].T, -1) @ v[0, :, s])
    want = torch.cat(outs, -1) @ p["wo"]["kernel"].reshape(d, d) + p["wo"]["bias"]
    assert torch.allclose(y[0], want, atol=1e-5)


def test_jumbo_residual_is_post_norm():
    cfg = O.tiny()
    p = O.init_params(cfg, 0, randomize_all=True)["model"]
    x = torch.randn(2, cfg.keep + 3, cfg.dim)
    lp = p["layer_0"]
    y = O.jumbo_layer(x, lp, p["jumbo_mlp"], cfg.heads, 3)
    x1 = x + O.attention(O.layer_norm(x, lp["norm1"]["scale"], lp["norm1"]["bias"]), lp["attn"], cfg.heads)
    c = O.layer_norm(x1[:, :3].reshape(2, -1), lp["norm3"]["scale"], lp["norm3"]["bias"])
    want_c = c + O.feed_forward(c, p["jumbo_mlp"])
    assert torch.allclose(y[:, :3].reshape(2, -1), want_c, atol=1e-5)


def test_lr_schedule_endpoints():
    cfg = O.vit_b16()
    peak = 1.5e-4 * 4096 / 256
    assert abs(O.lr_schedule(0, cfg) - 1e-6) < 1e-12
    assert abs(O.lr_schedule(cfg.warmup_steps, cfg) - peak) < 1e-12
    assert abs(O.lr_schedule(cfg.training_steps, cfg) - 1e-5) < 1e-12
    assert abs(O.lr_schedule(cfg.training_steps + 100, cfg) - 1e-5) < 1e-12
    half = cfg.warmup_steps // 2
    assert abs(O.lr_schedule(half, cfg) - (1e-6 + (peak - 1e-6) * half / cfg.warmup_steps)) < 1e-9


def test_adamw_first_step_and_decay_mask():
    cfg = O.vit_b16()
    p, g = torch.tensor([1.0, -2.0]), torch.tensor([0.5, -0.25])
    z = torch.zeros(2)
    new_p, m, v = O.adamw_leaf(p, g, z, z, 0, 1e-3, cfg, decay=False)
    # t=1: m_hat = g, v_hat = g^2 -> u = g / (|g| + eps) = sign(g)
    assert torch.allclose(new_p, p - 1e-3 * torch.sign(g), atol=1e-9)
    assert torch.allclose(m, 0.1 * g) and torch.allclose(v, 0.05 * g * g)
    new_p2, _, _ = O.adamw_leaf(p, g, z, z, 0, 1e-3, cfg, decay=True)
    assert torch.allclose(new_p2, new_p - 1e-3 * 0.05 * p, atol=1e-9)


def test_weight_decay_only_on_kernel_leaves():
    cfg = O.tiny()
    st = O.create_state(cfg, randomize_all=True)
    images, noise = O.synthetic_batch(cfg, 2)
    zero = O.tree_map(torch.zeros_like, st.params)
    st2, _, _ = O.training_step(st, images, noise, grads_override=zero)
    lr = O.lr_schedule(0, cfg)
    for path, new in O.tree_leaves(st2.params):
        old = dict(O.tree_leaves(st.params))[path]
        if path[-1] == "kernel":
            assert torch.allclose(new, old * (1 - lr * cfg.weight_decay), atol=1e-9)
        else:
            assert torch.equal(new, old)


def test_grads_match_finite_differences_f64():
    cfg = O.golden_tiny()
    st = O.create_state(cfg, dtype=torch.float64, randomize_all=True)
    images, noise = O.synthetic_batch(cfg, 2)
    loss, grads = O.loss_and_grads(st.params, images, noise, cfg)
    leaf = st.params["model"]["jumbo_mlp"]["w1"]["kernel"]
    g = grads["model"]["jumbo_mlp"]["w1"]["kernel"]
    eps = 1e-6
    for idx in [(0, 0), (5, 7), (20, 100)]:
        old = leaf[idx].item()
        leaf[idx] = old + eps
        lp = O.pretrain_forward(st.params, images, noise, cfg)["loss"].item()
        leaf[idx] = old - eps
        lm = O.pretrain_forward(st.params, images, noise, cfg)["loss"].item()
        leaf[idx] = old
        assert abs((lp - lm) / (2 * eps) - g[idx].item()) < 1e-6 * max(1.0, abs(g[idx].item()) * 1e3)


@pytest.mark.parametrize("name,norm_pix", [("tiny", False), ("tiny_normpix", True)])
def test_oracle_reproduces_committed_golden(name, norm_pix):
    z = np.load(GOLDEN)
    cfg = O.golden_tiny(norm_pix_loss=norm_pix)
    state = O.create_state(cfg, seed=0, dtype=torch.float64, randomize_all=True)
    # params come from the fixture, not from torch's RNG
    for path, leaf in O.tree_leaves(state.params):
        leaf.copy_(torch.from_numpy(z["tiny/params/" + "/".join(path)]))
    images = torch.from_numpy(z[f"{name}/images"])
    noise = z[f"{name}/noise"]
    collect = {}
    loss, grads = O.loss_and_grads(state.params, images, noise, cfg, collect)
    assert abs(float(loss) - float(z[f"{name}/loss"])) < 1e-10
    assert np.array_equal(collect["ids_restore"].numpy(), z[f"{name}/ids_restore"])
    assert np.array_equal(collect["mask"].numpy().astype(np.float32), z[f"{name}/mask"])
    for k in ("enc_in", "enc_out", "dec_in", "dec_out", "image_output"):
        assert np.allclose(collect[k].detach().numpy(), z[f"{name}/act/{k}"], atol=1e-5)
    for path, g in O.tree_leaves(grads):
        s = z[f"{name}/grad_stats/" + "/".join(path)]
        assert abs(g.sum().item() - s[0]) < 1e-9 + 1e-9 * abs(s[0])
        assert abs(g.norm().item() - s[1]) < 1e-9 + 1e-9 * abs(s[1])
    new_state, metrics, _ = O.training_step(state, images, noise)
    assert abs(metrics["learning_rate"] - float(z[f"{name}/lr"])) < 1e-15
    old = dict(O.tree_leaves(state.params))
    for path, p in O.tree_leaves(new_state.params):
        d = p - old[path]
        s = z[f"{name}/delta_stats/" + "/".join(path)]
        assert abs(d.norm().item() - s[1]) < 1e-12 + 1e-7 * abs(s[1])


def test_f32_oracle_close_to_f64_golden():
    z = np.load(GOLDEN)
    cfg = O.golden_tiny()
    state = O.create_state(cfg, seed=0, dtype=torch.float32, randomize_all=True)
    for path, leaf in O.tree_leaves(state.params):
        leaf.copy_(torch.from_numpy(z["tiny/params/" + "/".join(path)]).float())
    loss, grads = O.loss_and_grads(state.params, torch.from_numpy(z["tiny/images"]), z["tiny/noise"], cfg)
    assert abs(float(loss) - float(z["tiny/loss"])) < 1e-5
    for path, g in O.tree_leaves(grads):
        want = torch.from_numpy(z["tiny/grads/" + "/".join(path)])
        assert (g - want).norm() <= 1e-4 * want.norm() + 1e-7


# ------------------------------------------------------------------ fine-tune classification loss (8(f) rank 1)
def test_softmax_ce_known_answers():
    """Uniform logits give log C for any label distribution that sums to one; a dominant correct logit gives ~0."""
    C = 1000
    lab = np.eye(C, dtype=np.float32)[[3, 7]]
    assert np.allclose(O.softmax_cross_entropy(np.zeros((2, C), np.float32), lab), math.log(C), rtol=1e-6)
    lg = np.zeros((2, C), np.float32)
    lg[0, 3] = lg[1, 7] = 50.0
    assert np.all(O.softmax_cross_entropy(lg, lab) < 1e-6)


def test_smooth_labels_is_optax_rule():
    out = O.smooth_labels(np.eye(4, dtype=np.float32)[[1]], 0.1)
    assert np.allclose(out, [[0.025, 0.925, 0.025, 0.025]], atol=1e-7) and abs(out.sum() - 1) < 1e-6


def test_top_k_ties_go_to_the_lower_index():
    assert O.top_k_indices(np.array([[1.0, 3.0, 3.0, 0.0, 3.0, 2.0]], np.float32), 5).tolist() == [[1, 2, 4, 5, 0]]


def test_finetune_metrics_mixup_and_eval_modes():
    """acc counts a prediction inside the arg-max label SET (src/finetuning.py:99-106): with ratio 0.5 both mixed
    classes count; evaluation (det=True) ignores smoothing and mixing."""
    lg = np.zeros((2, 6), np.float32)
    lg[0, 4], lg[1, 0] = 2.0, 2.0
    m = O.finetune_metrics(lg, np.array([1, 4]), 6, 0.1, perm=np.array([1, 0]), ratio=0.5)
    assert m["acc1"].tolist() == [1.0, 0.0] and m["acc5"].tolist() == [1.0, 1.0]
    e = O.finetune_metrics(lg, np.array([1, 4]), 6, 0.1, perm=np.array([1, 0]), ratio=0.5, det=True)
    assert np.array_equal(e["labels"], np.eye(6, dtype=np.float32)[[1, 4]])
    g = O.softmax_ce_grad(lg, m["labels"])
    assert np.allclose(g.sum(-1), 0, atol=1e-7)      # labels sum to one -> every gradient row sums to zero


def test_cutmix_box_and_label_ratio():
    """random_bounding_box (src/utils.py:85-92): ratio 0.75 -> a box of half the side; the label weight of the pasted
    image is the box's share of the pixels; mixup with ratio 1 is the identity."""
    xm, ym, mask = O.random_bounding_box(0.75, 0.5, 0.5, 9, 9)
    assert xm.tolist() == [False, False, True, True, True, True, False, False, False]      # 0.25 <= x/8 < 0.75
    assert mask.shape == (1, 9, 9, 1) and (~mask).sum() == 16
    img = np.arange(2 * 9 * 9 * 3, dtype=np.float32).reshape(2, 9, 9, 3)
    lab = np.eye(2, dtype=np.float32)
    out, l2 = O.apply_cutmix(img, lab, 0.75, 0.5, 0.5, [1, 0])
    assert np.array_equal(out[0, 3, 3], img[1, 3, 3]) and np.array_equal(out[0, 0, 0], img[0, 0, 0])
    assert np.allclose(l2[0], [1 - 16 / 81, 16 / 81], atol=1e-6)
    same, l3 = O.apply_mixup(img, lab, 1.0, [1, 0])
    assert np.array_equal(same, img) and np.array_equal(l3, lab)


def test_layer_index_labels_and_scales():
    """get_layer_index_fn (src/utils.py:142-147) and the scale table of src/pretraining.py:235-242."""
    import dataclasses
    cfg = dataclasses.replace(O.vit_b16(), lr_decay=0.75)
    assert O.get_layer_index(("model", "embed", "wte", "kernel"), 12) == 0
    assert O.get_layer_index(("model", "layer_0", "attn", "wq", "kernel"), 12) == 1
    assert O.get_layer_index(("model", "layer_11", "norm3", "scale"), 12) == 12
    for path in (("model", "jumbo_mlp", "w1", "kernel"), ("model", "cls_tokens"), ("model", "norm", "scale"),
                 ("decoder_model", "dec_layer_3", "ff", "w2", "bias"), ("image_mask_embedding",), ("decoder_proj", "kernel")):
        assert O.get_layer_index(path, 12) == 12 and O.layerwise_scale(path, cfg) == 1.0
    assert O.layerwise_scale(("model", "embed", "wte", "bias"), cfg) == 0.75 ** 12
    assert O.layerwise_scale(("model", "layer_5", "ff", "w1", "kernel"), cfg) == 0.75 ** 6
    assert O.layerwise_scale(("model", "embed", "wte", "bias"), O.vit_b16()) == 1.0


def test_product_layer_index_matches_oracle():
    from jumbo_mae_tpu_b200.params import get_layer_index
    for path, _ in O.tree_leaves(O.init_params(O.tiny())):
        assert get_layer_index("/".join(path), 7) == O.get_layer_index(path, 7)


GOLDEN_CLASSIFY = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "tiny_classify.npz")


def _classify_from_golden(dtype):
    z = np.load(GOLDEN_CLASSIFY)
    cfg = O.golden_tiny(labels=int(z["logits"].shape[1]))
    params = O.init_params(cfg, seed=0, dtype=dtype, randomize_all=True)
    for path, leaf in O.tree_leaves(params):                  # params come from the fixture, not from torch's RNG
        leaf.copy_(torch.from_numpy(z["params/" + "/".join(path)]).to(dtype))
    images, perm, ratio = torch.from_numpy(z["images"]), z["perm"], float(z["ratio"])
    lab = O.mix_labels(O.smooth_labels(np.eye(cfg.labels, dtype=np.float32)[z["idx"]], float(z["smoothing"])), perm, ratio)
    norm = O.preprocess(images, dtype=dtype).numpy()
    r = norm.dtype.type(np.float32(ratio))
    x = r * norm + (1 - r) * norm[perm]
    return z, cfg, O.finetune_loss_and_grads(params, torch.as_tensor(x, dtype=dtype), lab, cfg), lab


def test_oracle_reproduces_committed_classification_golden():
    z, cfg, (loss, logits, grads), lab = _classify_from_golden(torch.float64)
    assert np.array_equal(lab, z["labels"])
    assert abs(float(loss) - float(z["loss"])) < 1e-10
    assert np.allclose(logits.numpy(), z["logits"], atol=1e-10)
    for path, g in O.tree_leaves(grads):
        s = z["grad_stats/" + "/".join(path)]
        assert abs(g.sum().item() - s[0]) < 1e-9 + 1e-9 * abs(s[0])
        assert abs(g.norm().item() - s[1]) < 1e-9 + 1e-9 * abs(s[1])


def test_f32_classification_oracle_close_to_f64_golden():
    z, cfg, (loss, logits, grads), _ = _classify_from_golden(torch.float32)
    assert abs(float(loss) - float(z["loss"])) < 1e-5
    assert np.allclose(logits.numpy(), z["logits"], atol=1e-4)
    for path, g in O.tree_leaves(grads):
        s = z["grad_stats/" + "/".join(path)]
        assert abs(g.double().norm().item() - s[1]) <= 1e-4 * s[1] + 1e-7


def test_lamb_leaf_trust_ratio_and_zero_norm_rule():
    """modified_lamb (src/utils.py:124-139): kernel leaves take u * ||p|| / ||u|| with u = adam + wd * p; a zero parameter
    (or update) norm gives ratio 1; other leaves are a plain Adam step without weight decay."""
    cfg = O.tiny()
    g = torch.Generator().manual_seed(0)
    p, grad = torch.randn(6, 5, generator=g), torch.randn(6, 5, generator=g) * 1e-2
    z = torch.zeros_like(p)
    new_p, m, v = O.lamb_leaf(p, grad, z, z, 0, 1e-2, cfg, decay=True)
    m_hat, v_hat = m / (1 - cfg.adam_b1), v / (1 - cfg.adam_b2)
    u = m_hat / (torch.sqrt(v_hat) + cfg.adam_eps) + cfg.weight_decay * p
    want = p - 1e-2 * u * (p.norm() / u.norm())
    assert torch.allclose(new_p, want, rtol=1e-6, atol=1e-8)
    assert torch.allclose((new_p - p).norm(), torch.tensor(1e-2) * p.norm(), rtol=1e-5)   # step length = lr * ||p||
    zero_p, _, _ = O.lamb_leaf(z, grad, z, z, 0, 1e-2, cfg, decay=True)                    # ||p|| = 0 -> ratio 1
    plain, _, _ = O.adamw_leaf(z, grad, z, z, 0, 1e-2, cfg, decay=True)
    assert torch.allclose(zero_p, plain)
    nodecay, _, _ = O.lamb_leaf(p, grad, z, z, 0, 1e-2, cfg, decay=False)
    adam, _, _ = O.adamw_leaf(p, grad, z, z, 0, 1e-2, cfg, decay=False)
    assert torch.allclose(nodecay, adam)


def test_per_image_masking_is_the_shared_rule_applied_row_by_row():
    """The per-image extension of random_masking (noise [B, L]): every image gets what the reference's rule
    (src/utils_mae.py:88-96) gives for its own noise row."""
    g = torch.Generator().manual_seed(1)
    B, L, D, keep = 3, 16, 4, 4
    x = torch.randn(B, L, D, generator=g)
    noise = np.random.default_rng(2).random((B, L), dtype=np.float32)
    kept, mask, ids_restore = O.random_masking(x, noise, keep)
    assert kept.shape == (B, keep, D) and mask.shape == (B, L) and ids_restore.shape == (B, L)
    for b in range(B):
        k1, m1, r1 = O.random_masking(x[b:b + 1], noise[b], keep)
        assert torch.equal(kept[b], k1[0]) and torch.equal(mask[b], m1[0]) and torch.equal(ids_restore[b], r1)
    # un-shuffling the kept tokens + mask tokens with ids_restore puts every kept token back at its own position
    full = torch.cat([kept, torch.zeros(B, L - keep, D)], dim=1)
    restored = O.index_sequence(full, ids_restore)
    assert torch.equal(restored[mask == 0], x[mask == 0])


def test_layerscale_scales_the_branches():
    """scale1/2/3 multiply the branch outputs (src/modeling.py:165-166,192-201): with all scales 0 a block is the identity
    on the patch tokens and LayerNorm-3 on the CLS tokens; the parameter tree carries them at 1e-4 (init.constant)."""
    cfg = O.tiny(layerscale=True, dec_layerscale=True)
    p = O.init_params(cfg, 0)
    lay = p["model"]["layer_0"]
    assert torch.all(lay["scale1"] == 1e-4) and lay["scale3"].shape == (cfg.num_cls_tokens * cfg.dim,)
    assert torch.all(p["decoder_model"]["dec_layer_0"]["scale2"] == 1e-4)
    g = torch.Generator().manual_seed(3)
    x = torch.randn(2, cfg.num_cls_tokens + 5, cfg.dim, generator=g)
    zeroed = {**lay, "scale1": torch.zeros(cfg.dim), "scale2": torch.zeros(cfg.dim),
              "scale3": torch.zeros(cfg.num_cls_tokens * cfg.dim)}
    y = O.jumbo_layer(x, zeroed, p["model"]["jumbo_mlp"], cfg.heads, cfg.num_cls_tokens)
    J = cfg.num_cls_tokens
    assert torch.equal(y[:, J:], x[:, J:])
    c = O.layer_norm(x[:, :J].reshape(2, -1), lay["norm3"]["scale"], lay["norm3"]["bias"])
    assert torch.allclose(y[:, :J].reshape(2, -1), c)
