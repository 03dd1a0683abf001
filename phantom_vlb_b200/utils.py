"""Host-side mirror of the reference's src/utils.py (same class names and call signatures), backed by libvlb_b200.

  get_hrf_weight                         : src/utils.py:14-37 -- nilearn's `compute_regressor(([0],[1],[1]), "glover",
      frame_times=[0, dt])[-1, 0]`: the Glover HRF (gamma pdfs, delay 6 / undershoot 12 / dispersion 0.9 / ratio 0.35, unit
      sum) convolved with a 1 s unit event on a 50x oversampled grid, read at dt by linear interpolation. nilearn is not
      needed: a boxcar convolved with h is a difference of two prefix sums of h, so one weight costs O(len(h)).
      `hrf_weights` maps it over arrays (the per-token Python loops of lazyloading.py:113-115, 135-137).
  HRFConvolveLayer, RidgeRegressionLayer : src/utils.py:40-73 -- callable, differentiable layers on the C ABI
      (vlb_hrf_convolve_fwd/bwd; vlb_gemm_bf16 + vlb_l2_prefill + vlb_colsum). Inside VLBLitModule.training_step both are
      fused into the head kernels of the engine; `litmodule.hrf_layer` / `.ridge_layer` are these same classes.
  LogValAccuracyCallback                 : src/utils.py:85-110 -- same hooks, but the per-parcel Pearson r is the
      single-pass CUDA kernel (vlb_pearson_*), row shards are merged with one all-reduce of the [5,P] moments, and
      the P correlations leave the GPU in ONE copy instead of P `self.log` syncs.
"""
import functools
import math

import numpy as np
import torch

try:                                    # Lightning is optional: absent in this image, present on a user's cluster
    from lightning.pytorch.callbacks import Callback
except Exception:                       # pragma: no cover
    class Callback:                     # minimal stand-in with the hook names the reference uses
        pass

from . import lib as _l
from . import ops

# nilearn.glm.first_level.hemodynamic_models: compute_regressor defaults and glover_hrf parameters
_OVERSAMPLING, _MIN_ONSET, _TIME_LENGTH = 50, -24.0, 32.0
_DELAY, _UNDERSHOOT, _DISPERSION, _RATIO = 6.0, 12.0, 0.9, 0.35


def _gamma_pdf(x, a, loc, scale):
    """scipy.stats.gamma.pdf(x, a, loc=loc, scale=scale) for a > 1 (0 for x <= loc)."""
    y = (np.asarray(x, dtype=np.float64) - loc) / scale
    out = np.zeros_like(y)
    pos = y > 0
    yp = y[pos]
    out[pos] = np.exp((a - 1.0) * np.log(yp) - yp - math.lgamma(a)) / scale
    return out


@functools.lru_cache(maxsize=65536)
def _glover_boxcar_weight(time_diff: float) -> float:
    dt_frame = float(time_diff)
    if not dt_frame > 0.0:
        return 0.0          # a token at or after the target time cannot have caused it (nilearn's frame grid is undefined here)
    # HRF kernel sampled every t_r / oversampling seconds for 32 s, unit sum (t_r = time_diff: two frame times)
    dt = dt_frame / _OVERSAMPLING
    ts = np.linspace(0, _TIME_LENGTH, int(np.rint(_TIME_LENGTH / dt)))
    h = _gamma_pdf(ts, _DELAY / _DISPERSION, dt, _DISPERSION) - _RATIO * _gamma_pdf(ts, _UNDERSHOOT / _DISPERSION, dt, _DISPERSION)
    h /= h.sum()
    # oversampled time grid of the event regressor: from min_onset before frame 0 to one frame past the last
    n_hr = int(np.rint((1.0 / dt_frame) * (2.0 * dt_frame - _MIN_ONSET) * _OVERSAMPLING + 1))
    grid = np.linspace(_MIN_ONSET, 2.0 * dt_frame, n_hr)
    t_on = min(int(np.searchsorted(grid, 0.0)), n_hr - 1)          # event: onset 0 s, duration 1 s, amplitude 1
    t_off = min(int(np.searchsorted(grid, 1.0)), n_hr - 1)
    if t_off < n_hr - 1 and t_off == t_on:
        t_off += 1
    # (boxcar * h)[k] = sum_{j=t_on}^{t_off-1} h[k-j] = H[k-t_on] - H[k-t_off],  H = prefix sums of h
    H = np.concatenate([[0.0], np.cumsum(h)])

    def conv(k):
        hi = min(max(k - t_on + 1, 0), h.size)
        lo = min(max(k - t_off + 1, 0), h.size)
        return H[hi] - H[lo]

    k1 = int(np.searchsorted(grid, dt_frame))                      # linear interpolation at the frame time
    k1 = min(max(k1, 1), n_hr - 1)
    k0 = k1 - 1
    f = (dt_frame - grid[k0]) / (grid[k1] - grid[k0])
    return float(conv(k0) + f * (conv(k1) - conv(k0)))


def get_hrf_weight(time_diff: float) -> float:
    """HRF weight of an input token `time_diff` seconds before the middle of the target TR (src/utils.py:14-37)."""
    return _glover_boxcar_weight(float(time_diff))


def hrf_weights(time_diffs):
    """get_hrf_weight over an array of lags (seconds) -> float64 array of the same shape (lazyloading.py:113-115, 135-137)."""
    t = np.asarray(time_diffs, dtype=np.float64)
    return np.array([_glover_boxcar_weight(float(x)) for x in t.ravel()], dtype=np.float64).reshape(t.shape)


# ------------------------------------------------------------------------------------------------ layers
def _cuda_bf16(t):
    if not t.is_cuda:
        raise RuntimeError("libvlb_b200 layers need CUDA tensors (there is no CPU fallback)")
    return t.detach().to(torch.bfloat16).contiguous()


class _HRFConvolveFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, embeddings, hrf_weights):
        B, S, H = embeddings.shape
        emb, w = _cuda_bf16(embeddings), hrf_weights.detach().to(device=embeddings.device, dtype=torch.float32).contiguous()
        out = torch.empty((B, H), device=emb.device, dtype=torch.float32)
        _l.call("vlb_hrf_convolve_fwd", _l.ptr(emb), _l.ptr(w), _l.ptr(out), B, S, H, _l.cur_stream())
        ctx.save_for_backward(emb, w)
        ctx.dtypes = (embeddings.dtype, hrf_weights.dtype)
        return out.to(embeddings.dtype)

    @staticmethod
    def backward(ctx, g):
        emb, w = ctx.saved_tensors
        B, S, H = emb.shape
        g = g.detach().to(torch.float32).contiguous()
        demb = torch.empty_like(emb) if ctx.needs_input_grad[0] else None
        dw = torch.empty((B, S), device=emb.device, dtype=torch.float32) if ctx.needs_input_grad[1] else None
        if demb is None and dw is None:
            return None, None
        _l.call("vlb_hrf_convolve_bwd", _l.ptr(emb), _l.ptr(w), _l.ptr(g), _l.ptr(demb), _l.ptr(dw), B, S, H, _l.cur_stream())
        return (demb.to(ctx.dtypes[0]) if demb is not None else None, dw.to(ctx.dtypes[1]) if dw is not None else None)


class HRFConvolveLayer(torch.nn.Module):
    """Weighted sum of token embeddings by HRF weight: einsum('bse,bs->be') (src/utils.py:40-56)."""

    def forward(self, embeddings, hrf_weights):
        return _HRFConvolveFn.apply(embeddings, hrf_weights)


class _RidgeFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, weight, bias, l2_lambda, add_reg):
        xb, wb, bb = _cuda_bf16(x.reshape(-1, x.shape[-1])), _cuda_bf16(weight), _cuda_bf16(bias)
        out = ops.gemm(xb, wb, bias=bb)
        acc = torch.zeros(2, device=xb.device, dtype=torch.float32)
        if add_reg:                                    # acc[1] = sum W^2 (vlb_l2_prefill without the gradient write)
            n = wb.numel()
            assert n % 8 == 0, "ridge weight element count must be a multiple of 8"
            _l.call("vlb_l2_prefill", _l.ptr(wb), None, _l.ptr(acc), n, float(l2_lambda), 0, _l.cur_stream())
        ctx.save_for_backward(xb, wb)
        ctx.l2_lambda, ctx.shape, ctx.dtypes = float(l2_lambda), x.shape, (x.dtype, weight.dtype, bias.dtype)
        return out.reshape(*x.shape[:-1], wb.shape[0]).to(x.dtype), (l2_lambda * acc[1]).to(weight.dtype)

    @staticmethod
    def backward(ctx, g_out, g_l2):
        xb, wb = ctx.saved_tensors
        P, H = wb.shape
        gy = _cuda_bf16(g_out.reshape(-1, P))
        dx = dw = db = None
        if ctx.needs_input_grad[0]:
            dx = ops.gemm(gy, wb, b_mn=True).reshape(ctx.shape).to(ctx.dtypes[0])
        if ctx.needs_input_grad[1]:
            dw32 = torch.zeros((P, H), device=wb.device, dtype=torch.float32)
            ops.gemm(gy, xb, dw32, a_mn=True, b_mn=True, out_f32=True, accumulate=True)
            if g_l2 is not None:
                dw32 += (2.0 * ctx.l2_lambda * g_l2.float()) * wb.float()
            dw = dw32.to(ctx.dtypes[1])
        if ctx.needs_input_grad[2]:
            db32 = torch.zeros((P,), device=wb.device, dtype=torch.float32)
            _l.call("vlb_colsum", _l.ptr(gy), _l.ptr(db32), gy.shape[0], P, _l.cur_stream())
            db = db32.to(ctx.dtypes[2])
        return dx, dw, db, None, None


class RidgeRegressionLayer(torch.nn.Module):
    """Linear readout + l2_lambda * ||W||^2 (src/utils.py:59-73); bf16 tensor-core arithmetic with fp32 accumulation, as the
    reference's layer under `precision: bf16-mixed`. Stand-alone it owns `linear.weight` / `linear.bias` like the reference;
    VLBLitModule binds it to the readout slice of its flat trainable buffer instead (`bind`)."""

    def __init__(self, input_dim, output_dim, l2_lambda=0.01, own_parameters=True, **kwargs):
        super().__init__()
        self.input_dim, self.output_dim, self.l2_lambda = input_dim, output_dim, l2_lambda
        self._bound = None
        if own_parameters:
            self.linear = torch.nn.Linear(input_dim, output_dim, bias=True, **kwargs)

    def bind(self, weight_fn, bias_fn):
        """Use tensors owned elsewhere (callables returning the current [P, H] weight and [P] bias)."""
        self._bound = (weight_fn, bias_fn)

    def forward(self, x, add_regularization=True):
        if self._bound is not None:
            w, b = self._bound[0](), self._bound[1]()
        else:
            w, b = self.linear.weight, self.linear.bias
        out, l2 = _RidgeFn.apply(x, w, b, self.l2_lambda, bool(add_regularization))
        return (out, l2) if add_regularization else out


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
