"""BASELINE configs[3] and configs[4] at full size on one GPU (parity-test cases with timing, not bench.py lines).

  python tools/bench_configs.py c4 [B] [P]   frozen 7B backbone, 16-frame clips: feature extraction clips/s, then the
                                             head-only step from cached features at P voxels (default 250 000)
  python tools/bench_configs.py c5 [N] [P]   per-voxel Pearson r over an [N, P] held-out season (fp32 and bf16 inputs)
Each line is JSON: measured time (CUDA events, warm), algorithmic bytes and the fraction of measured HBM peak.
"""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from bench import host_batches, peaks


def timeit(fn, warm=3, it=10):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(it):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / it


def c4(B=4, P=250000):
    from phantom_vlb_b200.engine import Engine
    from phantom_vlb_b200.model import ModelConfig, Trainables, frozen_random
    dev = torch.device("cuda", 0)
    _, hbm, _ = peaks()
    mc = ModelConfig(num_frames=16, num_target=P, use_lora=False)
    W = frozen_random(mc, dev, 1234)
    T = Trainables(mc, dev)
    T.init_default(1234)
    eng = Engine(mc, W, T, device=dev)
    host = host_batches(mc, B, 1, seed=1234)[0]
    batch = {k: v.to(dev) for k, v in host.items()}
    ms = timeit(lambda: eng.extract_features(batch), warm=2, it=4)
    tf = 39.0 * B                                   # 39.0 TF/clip frozen forward at T=16
    print(json.dumps(dict(kind="c4_extract_features", B=B, frames=16, ms=round(ms, 2), clips_per_s=round(B / ms * 1e3, 3),
                          tflops=round(tf / ms * 1e3, 1))), flush=True)
    # head-only training from cached features: N samples per step (the all-gathered global batch of an 8-GPU run)
    N = 64
    g = torch.Generator(device=dev).manual_seed(7)
    u = torch.randn((N, mc.hidden), device=dev, generator=g)
    sw = torch.rand((N,), device=dev, generator=g) + 0.5
    y = torch.randn((N, P), device=dev, generator=g)

    def step():
        eng.cached_loss_and_grads(u, sw, y)
        eng.optimizer_step(1e-4)

    ms = timeit(step, warm=2, it=5)
    H = mc.hidden
    alg = 3 * 2 * H * P                              # read W, read W for the penalty, write dW (bf16-sized)
    adam = 28 * (H * P)                              # fp32 p, g, m, v read + p, m, v write
    moved = (2 + 2 + 4 + 8 + 2) * H * P + adam       # what this build moves: W x3 (readout, penalty, d(hrf)), fp32 dW prefill+accumulate
    print(json.dumps(dict(kind="c4_cached_head_step", N=N, P=P, ms=round(ms, 3), loss=float(eng.buf("hd_loss", (3,), torch.float32)[0]),
                          alg_GB=round((alg + adam) / 1e9, 2), moved_GB_est=round(moved / 1e9, 2),
                          achieved_GBs=round((alg + adam) / ms / 1e6, 1), hbm_peak=hbm,
                          frac=round((alg + adam) / ms / 1e6 / hbm, 3))), flush=True)
    # parity at full size on a sample of voxels (fp32 torch on the same bf16-rounded operands). The wide-head path never
    # materialises dW, so the check is on the AdamW update itself: one more step from a snapshot of (W, m, v).
    idx = torch.randint(0, P, (512,), device=dev, generator=g)
    eng.cached_loss_and_grads(u, sw, y)
    hs = eng.head_saved
    pred_ref = hs["hrf"].float() @ T.w("ridge_w")[idx].float().t() + T.w("ridge_b")[idx].float()
    err = float((hs["pred"][:, idx].float() - pred_ref).abs().max())
    w0, m0, v0 = T.p("ridge_w")[idx].clone(), T.view(T.m, "ridge_w")[idx].clone(), T.view(T.v, "ridge_w")[idx].clone()
    gw = hs["dpred"][:, idx].float().t() @ hs["hrf"].float() + 2 * mc.l2_lambda * T.w("ridge_w")[idx].float()
    eng.optimizer_step(1e-4)
    clip = min(1.0, 1.0 / (float(T.sumsq) ** 0.5 + 1e-6))
    t = T.step_count
    gw = gw * clip
    m1 = 0.9 * m0 + 0.1 * gw
    v1 = 0.999 * v0 + 0.001 * gw * gw
    w1 = w0 * (1 - 1e-4 * 1e-2) - (1e-4 / (1 - 0.9 ** t)) * m1 / (v1.sqrt() / (1 - 0.999 ** t) ** 0.5 + 1e-8)
    uerr = float((T.p("ridge_w")[idx] - w1).abs().max() / (w1 - w0).abs().max())
    print(json.dumps(dict(kind="c4_full_size_parity", fused_head_opt=bool(eng.fused_head_opt), pred_max_abs_err=err,
                          adamw_update_rel_err=uerr, grad_norm=float(T.sumsq) ** 0.5)), flush=True)
    assert err < 3e-2 and uerr < 1e-3


def c5(N=5632, P=250000):
    from phantom_vlb_b200 import ops
    dev = torch.device("cuda", 0)
    _, hbm, _ = peaks()
    g = torch.Generator(device=dev).manual_seed(3)
    for dt in (torch.float32, torch.bfloat16):
        x = torch.randn((N, P), device=dev, generator=g).to(dt)
        y = (0.5 * x.float() + torch.randn((N, P), device=dev, generator=g)).to(dt)
        stats = torch.zeros((5, P), device=dev, dtype=torch.float64)

        def run():
            stats.zero_()
            ops.pearson(x, y, stats=stats, finalize=True)

        ms = timeit(run, warm=2, it=5)
        nbytes = 2 * N * P * x.element_size()
        r, _ = ops.pearson(x, y)
        cols = torch.arange(0, P, max(1, P // 64), device=dev)[:64]
        xs, ys = x[:, cols].double(), y[:, cols].double()
        xs, ys = xs - xs.mean(0), ys - ys.mean(0)
        r_ref = (xs * ys).sum(0) / (xs.pow(2).sum(0) * ys.pow(2).sum(0)).sqrt()
        err = float((r[cols].double() - r_ref).abs().max())
        print(json.dumps(dict(kind="c5_pearson", N=N, P=P, dtype=str(dt).split(".")[-1], ms=round(ms, 4),
                              alg_GB=round(nbytes / 1e9, 3), achieved_GBs=round(nbytes / ms / 1e6, 1), hbm_peak=hbm,
                              frac=round(nbytes / ms / 1e6 / hbm, 3), max_abs_err_vs_fp64=err)), flush=True)
        assert err < 1e-4
        del x, y


if __name__ == "__main__":
    which = sys.argv[1] if len(sys.argv) > 1 else "c5"
    args = [int(a) for a in sys.argv[2:]]
    {"c4": c4, "c5": c5}[which](*args)
