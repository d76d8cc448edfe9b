"""Host-side evaluation loop (src/main_pretrain.py:37-45, src/utils.py:34-50): CPU only."""
import numpy as np
import torch

from jumbo_mae_tpu_b200.evaluate import AverageMeter, evaluate


def test_average_meter_mean_latest_and_clear():
    m = AverageMeter(use_latest=["learning_rate"])
    m.update(loss=1.0, learning_rate=0.1)
    m.update(loss=torch.tensor([3.0]), learning_rate=0.2)
    out = m.summary("train/")
    assert out == {"train/loss": 2.0, "train/learning_rate": 0.2}
    assert m.summary() == {}


def test_evaluate_weights_batches_by_their_valid_samples():
    """validation_step returns sums + num_samples; a padded last batch (fewer valid rows) must not be over-weighted."""
    batches = [(np.full(4, 2.0), 4), (np.full(4, 2.0), 4), (np.full(1, 5.0), 1)]      # (per-sample losses, valid rows)

    def step(state, batch):
        vals, n = batch
        return {"loss": torch.tensor(vals.sum()), "acc1": float(n), "num_samples": torch.tensor(float(n))}

    out = evaluate(None, batches, step)
    assert abs(out["val/loss"] - (8 * 2.0 + 5.0) / 9) < 1e-12
    assert out["val/acc1"] == 1.0 and "val/num_samples" not in out
