"""GPU parity tests of the classification (fine-tune) step, SURVEY.md 8(f) rank 1: FinetuneEngine against the CPU
oracle's vit_classify / finetune_loss_and_grads (src/finetuning.py:84-124, src/modeling.py:245-274) on the same
parameters, images, labels and Mixup / CutMix draws.  Tolerances (bf16 kernels vs f32 oracle) beside each assert."""
import numpy as np
import pytest
import torch

from oracle import jumbo_mae_oracle as O

pytestmark = pytest.mark.gpu


def _rel(a, b):
    return ((a - b).norm() / (b.norm() + 1e-12)).item()


def _cfgs():
    return {
        "tiny": (O.tiny(labels=16), 4),
        "tiny_wpe": (O.tiny(labels=16, posemb="learnable"), 4),
        "tiny_ls": (O.tiny(labels=16, layerscale=True), 4),       # --layerscale: scale1/2/3 on the JumboLayer branches
        "mid": (O.MAEConfig(layers=2, dim=256, heads=4, patch_size=16, image_size=64, labels=40, warmup_steps=10,
                            training_steps=100, train_batch_size=256), 8),
    }


def _setup(name):
    from jumbo_mae_tpu_b200.config import MAEConfig
    from jumbo_mae_tpu_b200.finetune_engine import FinetuneEngine
    from jumbo_mae_tpu_b200.params import ParamStore
    ocfg, B = _cfgs()[name]
    cfg = MAEConfig(**{f: getattr(ocfg, f) for f in MAEConfig.__dataclass_fields__ if hasattr(ocfg, f)})
    params = O.init_params(ocfg, seed=0, randomize_all=True)
    images, _ = O.synthetic_batch(ocfg, B, seed=1)
    store = ParamStore(cfg)
    store.load_tree(params)
    return ocfg, cfg, params, images, store, FinetuneEngine(cfg, store, B)


@pytest.mark.parametrize("name,mode", [("tiny", "plain"), ("tiny", "mixup"), ("tiny", "cutmix"), ("mid", "mixup"),
                                       ("tiny_wpe", "plain"), ("tiny_ls", "plain")])
def test_classification_step_matches_oracle(name, mode):
    from jumbo_mae_tpu_b200.params import tree_leaves
    ocfg, cfg, params, images, store, eng = _setup(name)
    B, C, S = images.shape[0], ocfg.labels, ocfg.image_size
    rng = np.random.default_rng(3)
    idx = rng.integers(0, C, B).astype(np.int32)
    perm = rng.permutation(B).astype(np.int32)
    norm = O.preprocess(images).numpy()
    onehot = O.smooth_labels(np.eye(C, dtype=np.float32)[idx], 0.1)
    kw = {}
    if mode == "mixup":
        x, lab = O.apply_mixup(norm, onehot, 0.35, perm)
        kw = dict(perm=torch.from_numpy(perm).cuda(), ratio=0.35)
    elif mode == "cutmix":
        xm, ym, mask = O.random_bounding_box(0.6, 0.4, 0.7, S, S)
        x, lab = O.apply_cutmix(norm, onehot, 0.6, 0.4, 0.7, perm)
        kw = dict(perm=torch.from_numpy(perm).cuda(), xmask=torch.from_numpy(xm.astype(np.uint8)).cuda(),
                  ymask=torch.from_numpy(ym.astype(np.uint8)).cuda(),
                  label_ratio=float(mask.mean((1, 2), dtype=np.float32).item()))
    else:
        x, lab = norm, onehot
    want_loss, want_logits, want = O.finetune_loss_and_grads(params, x, lab, ocfg)
    loss = eng.forward(images.cuda(), torch.from_numpy(idx).cuda(), label_smoothing=0.1, **kw)
    eng.backward()
    eng.layerscale_finalize()
    torch.cuda.synchronize()
    logits = eng.logits.cpu()
    assert _rel(logits, want_logits) <= 2e-2                                   # bf16 activations, f32 accumulate
    assert abs(loss.item() - want_loss.item()) <= 5e-3 * abs(want_loss.item())
    # metrics of the engine's own logits, exactly (top-k of the oracle's logits may differ by a bf16 near-tie)
    m = O.finetune_metrics(logits.numpy(), idx, C, 0.1, perm=perm if mode != "plain" else None,
                           ratio=kw.get("label_ratio", kw.get("ratio", 1.0)))
    assert np.array_equal(eng.acc1.cpu().numpy(), m["acc1"]) and np.array_equal(eng.acc5.cpu().numpy(), m["acc5"])
    assert np.allclose(eng.loss_rows.cpu().numpy(), m["loss"], rtol=1e-5, atol=1e-6)
    assert np.allclose(m["labels"], lab, atol=1e-7)                            # the label matrix the oracle loss saw
    got = dict(tree_leaves(store.tree(store.g)))
    ref = dict(O.tree_leaves(want))
    bad = []
    for path, w in ref.items():
        g = got[path].detach().cpu().reshape(w.shape)
        if path[-2:] == ("wk", "bias"):           # identically zero gradient (softmax shift invariance): noise only
            scale = ref[path[:-2] + ("wv", "bias")].norm().item()
            if not (g.norm().item() <= 3e-2 * scale and w.norm().item() <= 1e-3 * scale):
                bad.append(("/".join(path), g.norm().item(), scale))
            continue
        if w.norm().item() == 0.0:
            # only the CLS rows reach the head: the last layer's patch FFN (and its norm2) has an exactly zero gradient
            assert path[1] == f"layer_{ocfg.layers - 1}" and path[2] in ("ff", "norm2", "scale2"), path
            if g.norm().item() != 0.0:
                bad.append(("/".join(path), g.norm().item(), 0.0))
            continue
        r = _rel(g, w)
        cos = (g.flatten() @ w.flatten() / (g.norm() * w.norm() + 1e-20)).item()
        if not (r <= 3e-2 and cos >= 0.999):
            bad.append(("/".join(path), r, cos))
    assert not bad, bad


def test_classification_eval_mode_and_accumulation():
    """det=True: no smoothing, no mixing, no gradient buffer written; backward accumulates like the MAE engine."""
    ocfg, cfg, params, images, store, eng = _setup("tiny")
    idx = torch.tensor([1, 5, 9, 2], dtype=torch.int32, device="cuda")
    eng.dlogits.fill_(7.0)
    loss = eng.forward(images.cuda(), idx, want_grad=False)
    want = O.vit_classify(O.preprocess(images), params["model"], ocfg)
    m = O.finetune_metrics(eng.logits.cpu().numpy(), idx.cpu().numpy(), ocfg.labels, det=True)
    assert _rel(eng.logits.cpu(), want) <= 2e-2
    assert abs(loss.item() - m["loss"].mean()) <= 1e-5 * abs(loss.item())
    assert torch.all(eng.dlogits == 7.0)
    eng.forward(images.cuda(), idx, label_smoothing=0.1)
    eng.backward()
    g1 = store.g.clone()
    eng.forward(images.cuda(), idx, label_smoothing=0.1)
    eng.backward(zero_grads=False)
    assert (store.g - 2 * g1).norm().item() <= 4e-3 * g1.norm().item()


def test_classification_step_at_vit_b_width():
    """ViT-B/16 width at 224 px (T = 199 tokens, head_dim 64, 1000 classes), one layer, against the oracle."""
    from jumbo_mae_tpu_b200.config import MAEConfig
    from jumbo_mae_tpu_b200.finetune_engine import FinetuneEngine
    from jumbo_mae_tpu_b200.params import ParamStore
    ocfg = O.MAEConfig(layers=1, labels=1000, warmup_steps=10, training_steps=100)
    cfg = MAEConfig(**{f: getattr(ocfg, f) for f in MAEConfig.__dataclass_fields__ if hasattr(ocfg, f)})
    params = O.init_params(ocfg, seed=0, randomize_all=True)
    images, _ = O.synthetic_batch(ocfg, 3, seed=2)
    store = ParamStore(cfg)
    store.load_tree(params)
    eng = FinetuneEngine(cfg, store, 3)
    idx = np.array([3, 999, 500], dtype=np.int32)
    lab = O.smooth_labels(np.eye(1000, dtype=np.float32)[idx], 0.1)
    want_loss, want_logits, want = O.finetune_loss_and_grads(params, O.preprocess(images).numpy(), lab, ocfg)
    loss = eng.forward(images.cuda(), torch.from_numpy(idx).cuda(), label_smoothing=0.1)
    eng.backward()
    assert _rel(eng.logits.cpu(), want_logits) <= 2e-2
    assert abs(loss.item() - want_loss.item()) <= 5e-3 * abs(want_loss.item())
    gk = store.grad("model/head/Dense_0/kernel")[:, :1000].cpu()
    assert _rel(gk, want["model"]["head"]["Dense_0"]["kernel"]) <= 3e-2
    ge = store.grad("model/embed/wte/kernel").cpu().reshape(-1)
    assert _rel(ge, want["model"]["embed"]["wte"]["kernel"].reshape(-1)) <= 3e-2


def _ft_args(**kw):
    import argparse
    base = dict(layers=2, dim=64, heads=2, labels=16, patch_size=8, image_size=32, init_seed=0, mixup_seed=0,
                dropout_seed=0, noise_seed=0, optimizer="adamw", criterion="ce", mode="finetune", learning_rate=1e-3,
                weight_decay=0.05, adam_b1=0.9, adam_b2=0.999, adam_eps=1e-8, lr_decay=0.75, clip_grad=0.0,
                grad_accum=1, warmup_steps=2, training_steps=50, posemb="sincos2d", dropout=0.0, droppath=0.0,
                layerscale=False, label_smoothing=0.1, mixup=0.8, cutmix=1.0, pretrained_ckpt=None)
    base.update(kw)
    return argparse.Namespace(**base)


@pytest.mark.parametrize("box", [None, (0.4, 0.7)])
def test_finetuning_api_step_matches_oracle(box):
    """create_train_state / training_step / validation_step (reference names, src/finetuning.py:108-284) with the
    Mixup / CutMix draws given: loss against the oracle on the state's own parameters, optimizer bookkeeping,
    validation sums over the rows whose label is not -1."""
    from jumbo_mae_tpu_b200.finetuning import MixDraw, create_train_state, training_step, validation_step, lr_schedule
    from jumbo_mae_tpu_b200.params import tree_leaves
    state = create_train_state(_ft_args()).replicate()
    ocfg = O.tiny(labels=16)
    params = O.init_params(ocfg, seed=0, randomize_all=True)
    state.store.load_tree(params)
    images, _ = O.synthetic_batch(ocfg, 4, seed=5)
    idx = np.array([1, 7, 7, 15], dtype=np.int32)
    perm = np.array([2, 3, 1, 0], dtype=np.int32)
    norm = O.preprocess(images).numpy()
    lab0 = O.smooth_labels(np.eye(16, dtype=np.float32)[idx], 0.1)
    if box is None:
        x, lab = O.apply_mixup(norm, lab0, 0.3, perm)
    else:
        x, lab = O.apply_cutmix(norm, lab0, 0.3, box[0], box[1], perm)
    want_loss, _, _ = O.finetune_loss_and_grads(params, x, lab, ocfg)
    before = state.store.p.clone()
    state, m = training_step(state, (images, torch.from_numpy(idx)), mix=MixDraw(perm=perm, ratio=0.3, box=box))
    assert abs(m["loss"].item() - want_loss.item()) <= 5e-3 * abs(want_loss.item())
    assert state.step == 1 and state.micro_step == 0
    assert m["learning_rate"] == lr_schedule(0, state) and abs(m["learning_rate"] - 1e-6) < 1e-15   # lr(0) = init
    assert not torch.equal(state.store.p, before)
    # validation: det=True, one padded row
    vidx = np.array([1, -1, 7, 15], dtype=np.int32)
    out = validation_step(state, (images, torch.from_numpy(vidx)))
    logits = state.engine(4).logits.cpu().numpy()
    om = O.finetune_metrics(logits, np.where(vidx == -1, 0, vidx), 16, det=True)
    keep = vidx != -1
    assert out["num_samples"].item() == 3
    assert abs(out["loss"].item() - om["loss"][keep].sum()) <= 1e-5 * abs(out["loss"].item())
    assert out["acc1"].item() == om["acc1"][keep].sum() and out["acc5"].item() == om["acc5"][keep].sum()


def test_finetuning_loss_goes_down_and_draws_advance():
    """A few steps on one batch without mixing reduce the loss; the default Mixup / CutMix draws differ step to step."""
    from jumbo_mae_tpu_b200.finetuning import create_train_state, training_step
    state = create_train_state(_ft_args(mixup=0.0, cutmix=0.0, learning_rate=3e-3, warmup_steps=1, lr_decay=1.0)).replicate()
    ocfg = O.tiny(labels=16)
    images, _ = O.synthetic_batch(ocfg, 8, seed=9)
    labels = torch.arange(8, dtype=torch.int32)
    losses = []
    for _ in range(12):
        state, m = training_step(state, (images, labels))
        losses.append(m["loss"].item())
    assert losses[-1] < 0.7 * losses[0], losses
    s2 = create_train_state(_ft_args()).replicate()
    d1, d2 = s2.draw_mix(8), s2.draw_mix(8)
    assert d1 is not None and (d1.ratio != d2.ratio or not np.array_equal(d1.perm, d2.perm))
    assert sorted(d1.perm.tolist()) == list(range(8)) and 0.0 <= d1.ratio <= 1.0


def test_load_pretrained_params_takes_the_encoder_and_keeps_the_head(tmp_path):
    """--pretrained-ckpt (src/finetuning.py:215-217, src/utils.py:150-202): a parameter file written by the
    pre-training step initialises the classification tree's encoder; head untouched, decoder ignored."""
    from jumbo_mae_tpu_b200 import checkpoint
    from jumbo_mae_tpu_b200.finetuning import create_train_state
    from jumbo_mae_tpu_b200.params import ParamStore, tree_leaves
    from jumbo_mae_tpu_b200.config import MAEConfig
    ocfg = O.tiny()
    pre = ParamStore(MAEConfig(**{f: getattr(ocfg, f) for f in MAEConfig.__dataclass_fields__ if hasattr(ocfg, f)}))
    pre.load_tree(O.init_params(ocfg, seed=3, randomize_all=True))
    path = str(tmp_path / "pretrain-last.msgpack")
    with open(path, "wb") as f:
        f.write(checkpoint.msgpack_serialize(pre.tree()))
    fresh = create_train_state(_ft_args())
    head0 = fresh.store.f("model/head/Dense_0/kernel").clone()
    state = create_train_state(_ft_args(pretrained_ckpt=path))
    assert torch.equal(state.store.f("model/head/Dense_0/kernel"), head0)          # same init seed, not overwritten
    want = dict(tree_leaves(pre.tree()["model"]))
    got = dict(tree_leaves(state.params["model"]))
    assert set(want) == {k for k in got if k[0] != "head"}
    for k, w in want.items():
        assert torch.equal(got[k], w), k
    assert torch.equal(state.store.p16, state.store.p.to(torch.bfloat16))          # compute copy refreshed


def test_classification_step_against_committed_golden():
    """FinetuneEngine on the committed fixture tests/golden/tiny_classify.npz (f64 oracle, make_golden_finetune.py):
    inputs, parameters and expected outputs come from the file, not from an RNG stream."""
    import os
    from jumbo_mae_tpu_b200.config import MAEConfig
    from jumbo_mae_tpu_b200.finetune_engine import FinetuneEngine
    from jumbo_mae_tpu_b200.params import ParamStore, tree_leaves
    z = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "tiny_classify.npz"))
    ocfg = O.golden_tiny(labels=int(z["logits"].shape[1]))
    cfg = MAEConfig(**{f: getattr(ocfg, f) for f in MAEConfig.__dataclass_fields__ if hasattr(ocfg, f)})
    params = O.init_params(ocfg, seed=0, randomize_all=True)
    for path, leaf in O.tree_leaves(params):
        leaf.copy_(torch.from_numpy(z["params/" + "/".join(path)]).float())
    store = ParamStore(cfg)
    store.load_tree(params)
    eng = FinetuneEngine(cfg, store, z["images"].shape[0])
    loss = eng.forward(torch.from_numpy(z["images"]).cuda(), torch.from_numpy(z["idx"]).cuda(),
                       label_smoothing=float(z["smoothing"]), perm=torch.from_numpy(z["perm"]).cuda(),
                       ratio=float(z["ratio"]))
    eng.backward()
    assert _rel(eng.logits.cpu().double(), torch.from_numpy(z["logits"])) <= 2e-2
    assert abs(loss.item() - float(z["loss"])) <= 5e-3 * float(z["loss"])
    got = dict(tree_leaves(store.tree(store.g)))
    for key in [k for k in z.files if k.startswith("grads/")]:
        path = tuple(key.split("/")[1:])
        w = torch.from_numpy(z[key])
        assert _rel(got[path].detach().cpu().reshape(w.shape), w) <= 3e-2, key
    for key in [k for k in z.files if k.startswith("grad_stats/")]:
        path = tuple(key.split("/")[1:])
        if path[-2:] == ("wk", "bias") or z[key][1] == 0.0:
            continue
        assert abs(got[path].norm().item() - z[key][1]) <= 3e-2 * z[key][1], key


@pytest.mark.parametrize("name", ["tiny", "mid", "tiny_ls"])
def test_droppath_matches_oracle(name):
    """DropPath (src/modeling.py:181-183,192,199,201; 0.1 in config/ft.sh) with the keep bits as data: some images lose
    a branch entirely, the others see it scaled by 1 / (1 - rate).  Logits, loss and every gradient leaf, same
    tolerances as the deterministic step."""
    from jumbo_mae_tpu_b200.params import tree_leaves
    ocfg, cfg, params, images, store, eng = _setup(name)
    B, C, rate = images.shape[0], ocfg.labels, 0.25
    rng = np.random.default_rng(11)
    keep = rng.random((ocfg.layers, 3, B)) >= rate
    keep[0, 0, 0] = keep[0, 1, 1] = keep[-1, 2, 2 % B] = False           # every branch kind is dropped somewhere
    fac = keep.astype(np.float32) / np.float32(1.0 - rate)
    idx = rng.integers(0, C, B).astype(np.int32)
    lab = O.smooth_labels(np.eye(C, dtype=np.float32)[idx], 0.1)
    want_loss, want_logits, want = O.finetune_loss_and_grads(params, O.preprocess(images).numpy(), lab, ocfg,
                                                            drop=torch.from_numpy(fac))
    eng.set_droppath(torch.from_numpy(fac).cuda())
    loss = eng.forward(images.cuda(), torch.from_numpy(idx).cuda(), label_smoothing=0.1)
    eng.backward()
    eng.layerscale_finalize()            # (tiny_ls: DropPath and LayerScale on the same branches)
    torch.cuda.synchronize()
    assert _rel(eng.logits.cpu(), want_logits) <= 2e-2
    assert abs(loss.item() - want_loss.item()) <= 5e-3 * abs(want_loss.item())
    got = dict(tree_leaves(store.tree(store.g)))
    bad = []
    ref = dict(O.tree_leaves(want))
    for path, w in ref.items():
        g = got[path].detach().cpu().reshape(w.shape)
        if path[-2:] == ("wk", "bias") or w.norm().item() == 0.0:
            continue                                                     # covered by the deterministic test
        r = _rel(g, w)
        cos = (g.flatten() @ w.flatten() / (g.norm() * w.norm() + 1e-20)).item()
        if not (r <= 3e-2 and cos >= 0.999):
            bad.append(("/".join(path), r, cos))
    assert not bad, bad
    # and the deterministic path is untouched once the masks are removed
    eng.set_droppath(None)
    eng.forward(images.cuda(), torch.from_numpy(idx).cuda(), label_smoothing=0.1, want_grad=False)
    det = O.vit_classify(O.preprocess(images), params["model"], ocfg)
    assert _rel(eng.logits.cpu(), det) <= 2e-2


def test_finetune_training_step_with_droppath_draws():
    """training_step with --droppath: the keep bits come from the rank's dropout stream, advance every step, and a
    caller-supplied draw reproduces the oracle's loss."""
    import argparse
    from jumbo_mae_tpu_b200.finetuning import create_train_state, training_step
    args = argparse.Namespace(layers=2, dim=64, heads=2, patch_size=8, image_size=32, labels=16, weight_decay=0.05,
                              adam_b1=0.9, adam_b2=0.999, adam_eps=1e-8, learning_rate=1e-3, warmup_steps=2,
                              training_steps=10, label_smoothing=0.1, mixup=0.0, cutmix=0.0, droppath=0.5, init_seed=0,
                              dropout_seed=3)
    state = create_train_state(args).replicate()
    a, b = state.draw_drop_keep(4), state.draw_drop_keep(4)
    assert a.shape == (2, 3, 4) and a.dtype == bool and not np.array_equal(a, b)
    images = torch.randint(0, 256, (4, 3, 32, 32), dtype=torch.uint8, generator=torch.Generator().manual_seed(0))
    labels = torch.tensor([1, 2, 3, 4], dtype=torch.int32)
    keep = np.ones((2, 3, 4), dtype=bool)
    keep[1, :, 0] = False
    ocfg = O.tiny(labels=16)
    params = O.tree_map(lambda t: t.detach().cpu().clone(), state.params)
    lab = O.smooth_labels(np.eye(16, dtype=np.float32)[labels.numpy()], 0.1)
    want, _, _ = O.finetune_loss_and_grads(params, O.preprocess(images).numpy(), lab, ocfg,
                                           drop=torch.from_numpy(keep.astype(np.float32) / 0.5))
    state, m = training_step(state, (images, labels), drop_keep=keep)
    assert abs(m["loss"].item() - want.item()) <= 5e-3 * abs(want.item())
    state, m = training_step(state, (images, labels))                       # draws from the stream
    assert np.isfinite(m["loss"].item())
