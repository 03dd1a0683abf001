"""Generates tests/golden/*.npz by EXECUTING the reference's own files from /root/reference (build container only;
the GPU box has no /root/reference).  Third-party imports that are absent here (lightning, nilearn, torchmetrics, timm,
peft, h5py, decord ...) are stubbed; timm's RegStage/LayerNorm2d are provided by oracle/stc_timm.py, so the
STCConnector golden pins the reference's *glue* (einops rearranges, sampler, readout), not timm's arithmetic.

    python -m oracle.make_goldens            # rewrites tests/golden/
"""
import importlib.machinery
import importlib.util
import os
import sys
import types

import numpy as np
import torch

REF = "/root/reference"
OUT = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden")


def _stub(name, **attrs):
    m = types.ModuleType(name)
    m.__spec__ = importlib.machinery.ModuleSpec(name, None)
    for k, v in attrs.items():
        setattr(m, k, v)
    sys.modules[name] = m
    return m


def _load(path, name):
    spec = importlib.util.spec_from_file_location(name, path)
    mod = importlib.util.module_from_spec(spec)
    sys.modules[name] = mod
    spec.loader.exec_module(mod)
    return mod


def ref_utils():
    _stub("lightning"); _stub("lightning.pytorch"); _stub("lightning.pytorch.callbacks", Callback=object)
    _stub("nilearn"); _stub("nilearn.glm"); _stub("nilearn.glm.first_level", compute_regressor=None)
    _stub("torchmetrics", PearsonCorrCoef=None)
    return _load(os.path.join(REF, "src/utils.py"), "ref_src_utils")


def ref_projector():
    from oracle import stc_timm
    _stub("timm"); _stub("timm.models"); _stub("timm.models.regnet", RegStage=stc_timm.RegStage)
    _stub("timm.models.layers", LayerNorm=torch.nn.LayerNorm, LayerNorm2d=stc_timm.LayerNorm2d)
    import transformers
    if not hasattr(transformers, "TRANSFORMERS_CACHE"):      # removed in transformers 5 (reference pins 4.53.3)
        transformers.TRANSFORMERS_CACHE = "/tmp"
    return _load(os.path.join(REF, "VideoLLaMA2/videollama2/model/projector.py"), "ref_projector")


def ref_mm_utils_frame_sample():
    src = open(os.path.join(REF, "VideoLLaMA2/videollama2/mm_utils.py")).read()
    start = src.index("def frame_sample(")
    end = src.index("\ndef ", start + 10)
    ns = {"np": np, "NUM_FRAMES_PER_SECOND": 1}
    exec(src[start:end], ns)
    return ns["frame_sample"]


def ref_make_weight_mask():
    """Extract VLBLitModule.make_weight_mask (litmodule.py:178-203) and bind it to a minimal object."""
    src = open(os.path.join(REF, "src/litmodule/videollama2_vlb_litmodule.py")).read()
    start = src.index("    def make_weight_mask(")
    end = src.index("    def configure_model(", start)
    import textwrap
    ns = {"torch": torch}
    exec(textwrap.dedent(src[start:end]), ns)

    class Self:
        device = torch.device("cpu")

        class config:
            dtype = torch.bfloat16
    return lambda *a: ns["make_weight_mask"](Self(), *a)


def main():
    os.makedirs(OUT, exist_ok=True)
    torch.manual_seed(1234)
    # ---- src/utils.py head layers
    U = ref_utils()
    emb, w = torch.randn(3, 40, 32), torch.rand(3, 40)
    pooled = U.HRFConvolveLayer()(emb, w)
    ridge = U.RidgeRegressionLayer(32, 24, l2_lambda=0.01)
    out, l2 = ridge(pooled)
    np.savez(os.path.join(OUT, "head_utils.npz"), emb=emb.numpy(), w=w.numpy(), pooled=pooled.detach().numpy(),
             ridge_w=ridge.linear.weight.detach().numpy(), ridge_b=ridge.linear.bias.detach().numpy(),
             out=out.detach().numpy(), l2=float(l2))
    # ---- make_weight_mask
    mk = ref_make_weight_mask()
    g = torch.Generator().manual_seed(7)
    padvals = torch.tensor([[5, 9, 20], [0, 9, 58], [300, 9, 2]])
    visw, langw = torch.rand(3, 5, generator=g, dtype=torch.float64), torch.rand(3, 64, generator=g, dtype=torch.float64)
    wm = mk(padvals, visw, langw, 1204, 2048)
    np.savez(os.path.join(OUT, "weight_mask.npz"), padvals=padvals.numpy(), vis_w=visw.numpy(), lang_w=langw.numpy(),
             mask=wm.float().numpy())
    # ---- frame_sample / chunk indices known answers
    fs = ref_mm_utils_frame_sample()
    durations = [44, 89, 133, 135, 7, 12, 500]
    np.savez(os.path.join(OUT, "frame_sample.npz"), durations=np.array(durations),
             **{"d%d_n%d" % (d, n): fs(d, "uniform", n) for d in durations for n in (4, 8, 12)})
    # ---- STCConnector glue on tiny widths
    P = ref_projector()

    class Cfg:
        mm_hidden_size, hidden_size = 32, 64
    torch.manual_seed(99)
    stc = P.STCConnector(Cfg(), depth=2)
    x = torch.randn(2, 4, 16, 32)
    y = stc(x)
    sd = {k: v.detach().numpy() for k, v in stc.state_dict().items()}
    np.savez(os.path.join(OUT, "stc_connector.npz"), x=x.numpy(), y=y.detach().numpy(), **{"sd." + k: v for k, v in sd.items()})
    print("goldens written to", OUT, sorted(os.listdir(OUT)))


if __name__ == "__main__":
    main()
