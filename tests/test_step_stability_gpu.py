"""Full-size (7B, 8-frame clips) repeatability: the forward has no atomics, so repeated validation steps on one batch must be
bit-identical, and a few consecutive train steps must stay finite and decrease the loss. Small-shape parity tests cannot see
ordering bugs that need hundreds of co-resident CTAs (r01d: intermittent NaN from an mbarrier phase aliasing)."""
import math

import pytest
import torch

pytestmark = pytest.mark.gpu


def _module_and_batch(B):
    from bench import host_batches, lit_config
    from phantom_vlb_b200.litmodule import VLBLitModule
    m = VLBLitModule(lit_config())
    m.configure_model(device=torch.device("cuda", 0))
    batch = {k: v.cuda() for k, v in host_batches(m.model_config, B, 1, 1234)[0].items()}
    return m, batch


def test_full_size_forward_is_bit_repeatable_and_training_stays_finite():
    m, batch = _module_and_batch(2)
    ref = m.validation_step(batch)["brain_preds"].float().clone()
    assert not torch.isnan(ref).any()
    for i in range(12):
        p = m.validation_step(batch)["brain_preds"].float()
        assert torch.equal(p, ref), "validation forward %d differs from the first" % i
    losses = [float(m.engine.train_step(batch, 1e-4)[0]) for _ in range(8)]
    assert all(math.isfinite(x) for x in losses), losses
    assert losses[-1] < losses[0], losses
