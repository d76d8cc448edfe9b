"""The reference-named Python boundary on the GPU: the helpers of ``utils_mae`` and the callable
modules ``ViT`` / ``MAEDecoder`` / ``PretrainModule`` against the oracle's functions of the same name, with the
reference's argument lists (src/utils_mae.py:51-102, src/modeling.py:245,286, src/pretraining.py:87).

Tolerances: index / mask work exact; f32 elementwise work rel 1e-5; module outputs as in tests/test_step_gpu.py
(bf16 operands, f32 accumulation): rel-L2 <= 1e-2 encoder, <= 3e-2 decoder, loss rel <= 5e-3."""
import numpy as np
import pytest
import torch

from oracle import jumbo_mae_oracle as O

pytestmark = pytest.mark.gpu


def _rel(a, b):
    return ((a - b).norm() / (b.norm() + 1e-12)).item()


@pytest.mark.parametrize("B,L,D,keep", [(3, 196, 64, 49), (2, 16, 8, 4), (1, 5, 4, 1)])
def test_random_masking_and_index_sequence(B, L, D, keep):
    from jumbo_mae_tpu_b200.utils_mae import index_sequence, random_masking
    g = torch.Generator().manual_seed(L)
    x = torch.randn(B, L, D, generator=g)
    noise = np.random.default_rng(L).random(L, dtype=np.float32)
    noise[1 % L] = noise[0]                                          # a tie: stable argsort keeps index order
    kept, mask, ids_restore = random_masking(x.cuda(), torch.from_numpy(noise).cuda(), keep)
    wk, wm, wi = O.random_masking(x, noise, keep)
    assert torch.equal(kept.cpu(), wk) and torch.equal(mask.cpu(), wm)
    assert np.array_equal(ids_restore.cpu().numpy(), wi.numpy())
    ids = torch.from_numpy(np.random.default_rng(1).permutation(L).astype(np.int32))
    assert torch.equal(index_sequence(x.cuda(), ids.cuda()).cpu(), O.index_sequence(x, ids))


def test_extract_patches_matches_oracle():
    from jumbo_mae_tpu_b200.utils_mae import extract_patches
    images = torch.randint(0, 256, (2, 3, 64, 64), dtype=torch.uint8, generator=torch.Generator().manual_seed(0))
    got = extract_patches(images.cuda(), 16).cpu()
    want = O.extract_patches(O.preprocess(images), 16)
    assert torch.equal(got, want)                                    # the same IEEE operations in the same order


@pytest.mark.parametrize("with_valid", [False, True])
def test_patch_mse_loss_reference_signature(with_valid):
    """patch_mse_loss(patch_output, patch_target, valid=None), src/utils_mae.py:51: per-image valid masks allowed."""
    from jumbo_mae_tpu_b200.utils_mae import patch_mse_loss
    g = torch.Generator().manual_seed(3)
    out, tgt = torch.randn(4, 49, 192, generator=g), torch.randn(4, 49, 192, generator=g)
    valid = (torch.rand(4, 49, generator=g) > 0.4).float() if with_valid else None
    got = patch_mse_loss(out.cuda(), tgt.cuda(), None if valid is None else valid.cuda()).item()
    want = O.patch_mse_loss(out, tgt, valid).item()
    assert abs(got - want) <= 1e-5 * abs(want)


def test_patch_mse_loss_from_images_equals_the_two_step_form():
    from jumbo_mae_tpu_b200.utils_mae import extract_patches, patch_mse_loss, patch_mse_loss_from_images
    g = torch.Generator().manual_seed(4)
    images = torch.randint(0, 256, (3, 3, 32, 32), dtype=torch.uint8, generator=g).cuda()
    out = torch.randn(3, 16, 192, generator=g).cuda()
    valid = (torch.rand(16, generator=g) > 0.5).float().cuda()
    fused = patch_mse_loss_from_images(out, images, valid, 8).item()
    two_step = patch_mse_loss(out, extract_patches(images, 8), valid[None, :].expand(3, 16)).item()
    assert abs(fused - two_step) <= 1e-5 * abs(two_step)


def _bound_modules(ocfg, B):
    from jumbo_mae_tpu_b200.config import MAEConfig
    from jumbo_mae_tpu_b200.modeling import MAEDecoder, ViT
    from jumbo_mae_tpu_b200.params import ParamStore
    from jumbo_mae_tpu_b200.pretraining import PretrainModule
    cfg = MAEConfig(**{f: getattr(ocfg, f) for f in MAEConfig.__dataclass_fields__ if hasattr(ocfg, f)})
    state = O.create_state(ocfg, seed=0, randomize_all=True)
    store = ParamStore(cfg)
    store.load_tree(state.params)
    images, noise = O.synthetic_batch(ocfg, B, seed=1, noise_seed=1001)
    vit = ViT(layers=ocfg.layers, dim=ocfg.dim, heads=ocfg.heads, labels=None, patch_size=ocfg.patch_size,
              image_size=ocfg.image_size, image_mask_ratio=ocfg.image_mask_ratio)
    dec = MAEDecoder(dec_layers=ocfg.dec_layers, dec_dim=ocfg.dec_dim, dec_heads=ocfg.dec_heads,
                     patch_size=ocfg.patch_size, image_size=ocfg.image_size)
    return state, store, images, noise, vit, dec, PretrainModule(vit, dec, ocfg.image_size, ocfg.norm_pix_loss)


def test_vit_call_mae_branch():
    """ViT.__call__(x, det=True) (src/modeling.py:245): (x, mask, ids_restore)."""
    ocfg = O.tiny()
    state, store, images, noise, vit, _, _ = _bound_modules(ocfg, 4)
    x, mask, ids_restore = vit.bind(store, rngs={"noise": noise})(images.cuda(), det=True)
    wx, wmask, wids = O.vit_forward(O.preprocess(images), state.params["model"], ocfg, noise)
    assert np.array_equal(ids_restore.cpu().numpy(), wids.numpy()) and torch.equal(mask.cpu(), wmask)
    assert _rel(x.cpu(), wx) <= 1e-2
    with pytest.raises(Exception):
        vit(images.cuda())                                           # unbound: no parameters, no noise


def test_mae_decoder_call():
    """MAEDecoder.__call__(x, det=True) (src/modeling.py:286): wpe on the patch tokens, blocks, dec_norm."""
    ocfg = O.tiny()
    state, store, _, _, _, dec, _ = _bound_modules(ocfg, 3)
    x = torch.randn(3, ocfg.num_cls_tokens + ocfg.num_patches, ocfg.dec_dim, generator=torch.Generator().manual_seed(2))
    got = dec.bind(store)(x.cuda(), det=True).cpu()
    want = O.decoder_forward(x, state.params["decoder_model"], ocfg)
    assert _rel(got, want) <= 3e-2


@pytest.mark.parametrize("norm_pix", [False, True])
def test_pretrain_module_call(norm_pix):
    """PretrainModule.__call__(images, det=True) (src/pretraining.py:87) -> {"loss": scalar}."""
    ocfg = O.tiny(norm_pix_loss=norm_pix)
    state, store, images, noise, _, _, module = _bound_modules(ocfg, 4)
    got = module.bind(store, rngs={"noise": torch.from_numpy(noise)})(images, det=True)
    want = O.pretrain_forward(state.params, images, noise, ocfg)["loss"].item()
    assert set(got) == {"loss"} and abs(got["loss"].item() - want) <= 5e-3 * abs(want)
