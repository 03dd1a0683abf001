"""Host-side mirror of the reference's src/utils.py (same class names and call signatures), backed by libvlb_b200.

  HRFConvolveLayer, RidgeRegressionLayer : src/utils.py:40-73 -- in this build both live inside the fused head
      kernels of the engine; the classes are kept so `litmodule.hrf_layer` / `.ridge_layer` exist (checkpoint keys).
  LogValAccuracyCallback                 : src/utils.py:85-110 -- same hooks, but the per-parcel Pearson r is the
      single-pass CUDA kernel (vlb_pearson_*), row shards are merged with one all-reduce of the [5,P] moments, and
      the P correlations leave the GPU in ONE copy instead of P `self.log` syncs.
  get_hrf_weight                         : src/utils.py:14-37 (offline helper; nilearn's Glover HRF restated with
      scipy-free closed form -- only used to build synthetic HRF weights, never inside the step).
"""
import math

import numpy as np
import torch

try:                                    # Lightning is optional: absent in this image, present on a user's cluster
    from lightning.pytorch.callbacks import Callback
except Exception:                       # pragma: no cover
    class Callback:                     # minimal stand-in with the hook names the reference uses
        pass

from . import ops


def get_hrf_weight(time_diff: float) -> float:
    """Double-gamma (Glover) HRF sampled at `time_diff` seconds after a unit impulse, peak-normalised."""
    t = float(time_diff)
    if t <= 0:
        return 0.0
    h = (t ** 5 * math.exp(-t) / math.gamma(6.0)) - (t ** 15 * math.exp(-t) / math.gamma(16.0)) / 6.0
    return h / 0.1758


def hrf_weights(time_diffs):
    """Vectorised get_hrf_weight over an array of lags (seconds): one numpy expression instead of one nilearn
    compute_regressor call per visual / language token (lazyloading.py:113-115, 135-137 -- the per-token Python loop the
    reference runs for every sample). Returns float64 like the reference's weight arrays."""
    t = np.asarray(time_diffs, dtype=np.float64)
    tp = np.where(t > 0, t, 1.0)
    h = (tp ** 5 * np.exp(-tp) / math.gamma(6.0)) - (tp ** 15 * np.exp(-tp) / math.gamma(16.0)) / 6.0
    return np.where(t > 0, h / 0.1758, 0.0)


class HRFConvolveLayer(torch.nn.Module):
    """Weighted sum of token embeddings by HRF weight: einsum('bse,bs->be'). Fused into vlb_hrf_pool_fwd."""

    def forward(self, embeddings, hrf_weights):
        raise RuntimeError("HRFConvolveLayer is fused into the CUDA head (Engine.head_forward); it is not called alone")


class RidgeRegressionLayer(torch.nn.Module):
    """Linear readout + l2_lambda * ||W||^2. Parameters live in the engine's flat trainable buffer."""

    def __init__(self, input_dim, output_dim, l2_lambda=0.01, **kwargs):
        super().__init__()
        self.input_dim, self.output_dim, self.l2_lambda = input_dim, output_dim, l2_lambda

    def forward(self, x, add_regularization=True):
        raise RuntimeError("RidgeRegressionLayer is fused into the CUDA head (Engine.head_forward)")


def pearson_columns(preds, targets, group=None):
    """Per-column Pearson r on the GPU; row shards on different ranks are merged with one all-reduce of the moments."""
    from .dp import merge_moments_
    N, P = preds.shape
    _, stats = ops.pearson(preds.contiguous(), targets.contiguous(), finalize=False)
    stats, n = merge_moments_(stats, N, group)
    return ops.pearson_finalize(stats, n)


class LogValAccuracyCallback(Callback):
    def __init__(self, csv_dir=None):
        """csv_dir: when set, every validation epoch also appends one row to <csv_dir>/metrics.csv with the reference's
        column names, so make_acc_brainmaps.py (and postprocessing.make_brainmaps) can read it without Lightning's logger."""
        super().__init__()
        self.csv_writer = None
        if csv_dir is not None:
            from .postprocessing import MetricsCSVWriter
            self.csv_writer = MetricsCSVWriter(csv_dir)

    def on_validation_epoch_start(self, trainer, pl_module):
        self.val_actual = []
        self.val_pred = []
        self.val_loss = []

    def on_validation_batch_end(self, trainer, pl_module, outputs, batch, batch_idx, dataloader_idx=0):
        self.val_actual.append(outputs["brain_vals"])
        self.val_pred.append(outputs["brain_preds"])
        if "loss" in outputs:
            self.val_loss.append(outputs["loss"].detach().clone())

    def on_validation_epoch_end(self, trainer, pl_module):
        all_vals = torch.cat(self.val_actual, dim=0)
        all_preds = torch.cat(self.val_pred, dim=0)
        correlations = pearson_columns(all_preds, all_vals)          # nan_to_num is applied inside the kernel
        host = correlations.float().cpu().numpy()                    # one D2H copy for all parcels
        pl_module.log_dict({f"val_corr_ROI_{i:0{6}}": float(host[i]) for i in range(pl_module.config.num_target)})
        pl_module.log("val_corr_avg", float(np.mean(host)))
        self.last_correlations = host
        if getattr(self, "csv_writer", None) is not None:            # columnar metrics.csv (postprocessing.MetricsCSVWriter)
            loss = torch.stack([v.float().mean() for v in self.val_loss]).mean().item() if getattr(self, "val_loss", None) else float("nan")
            self.csv_writer.log_validation(getattr(trainer, "current_epoch", 0), getattr(trainer, "global_step", 0), loss, host)
