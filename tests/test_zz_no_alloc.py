"""vlb_hrf_pool_fwd_static: the pooled head forward (final RMSNorm -> LN1 -> HRF pooling -> LN2 -> dropout) as ONE C-ABI call
that allocates nothing -- its partial[B, nch, H] scratch is a file-scope device array of the library (2 MiB), resolved with
cudaGetSymbolAddress once per device. No counterpart in the reference; added for callers that cannot call cudaMalloc.

Checked against the existing two-call form (vlb_hrf_pool_fwd + vlb_hrf_pool_finalize with a caller buffer): the same
kernels run on the same operands, so the results must be bit-identical (the other head tests allow 1e-2 relative; this is
stricter). Two calls in a row with different shapes, so the second one runs over the first one's leftovers in the shared
array. Collected last on purpose."""
import pytest
import torch

pytestmark = pytest.mark.gpu


def _inputs(B, S, H, seed):
    g = torch.Generator(device="cuda").manual_seed(seed)
    r = lambda *s: torch.randn(*s, device="cuda", generator=g)
    x = r(B, S, H).bfloat16()
    wm = torch.rand(B, S, device="cuda", generator=g)
    wm = torch.where(wm < 0.4, torch.zeros_like(wm), wm).float()          # rows with weight 0 are skipped by the kernel
    wm[:, : S // 8] = 0.0                                                 # a whole chunk of rows without any weight
    ws = [(1.0 + 0.1 * r(H)).bfloat16() for _ in range(3)]                # norm_w, ln1_w, ln2_w
    bs = [(0.1 * r(H)).bfloat16() for _ in range(2)]                      # ln1_b, ln2_b
    return x, wm, ws[0], ws[1], bs[0], ws[2], bs[1]


def _two_call(x, wm, nw, g1, b1, g2, b2, nch, drop, seed):
    from phantom_vlb_b200 import lib as _l
    B, S, H = x.shape
    partial = torch.full((B, nch, H), float("nan"), device="cuda")
    pooled, hrf = torch.empty(B, H, device="cuda", dtype=torch.bfloat16), torch.empty(B, H, device="cuda", dtype=torch.bfloat16)
    keep = torch.empty(B, H, device="cuda")
    _l.call("vlb_hrf_pool_fwd", _l.ptr(x), _l.ptr(nw), _l.ptr(g1), _l.ptr(b1), _l.ptr(wm), _l.ptr(partial), B, S, H, nch,
            1e-5, 1e-5, _l.cur_stream())
    _l.call("vlb_hrf_pool_finalize", _l.ptr(partial), _l.ptr(g2), _l.ptr(b2), _l.ptr(pooled), _l.ptr(hrf), _l.ptr(keep),
            B, H, nch, 1e-5, drop, seed, _l.cur_stream())
    return pooled, hrf, keep


def _static(x, wm, nw, g1, b1, g2, b2, nch, drop, seed, out):
    from phantom_vlb_b200 import lib as _l
    B, S, H = x.shape
    pooled, hrf, keep = out
    _l.call("vlb_hrf_pool_fwd_static", _l.ptr(x), _l.ptr(nw), _l.ptr(g1), _l.ptr(b1), _l.ptr(wm), _l.ptr(g2), _l.ptr(b2),
            _l.ptr(pooled), _l.ptr(hrf), _l.ptr(keep), B, S, H, nch, 1e-5, 1e-5, drop, seed, _l.cur_stream())


def test_static_scratch_pool_matches_two_call_form_on_two_calls_in_a_row():
    from phantom_vlb_b200 import lib as _l
    _l.call("vlb_check_device")
    assert _l.lib().vlb_hrf_pool_static_max_floats() == 512 * 1024
    # call 1 fills the whole array (8 * 16 * 4096 floats = the limit); call 2 is smaller and ragged (S not a multiple of nch)
    cases = [(8, 2048, 4096, 16, 0.1, 0x5EED01), (3, 1000, 1024, 7, 0.0, 0x5EED02)]
    data = [_inputs(B, S, H, 100 + i) for i, (B, S, H, _, _, _) in enumerate(cases)]
    want = [_two_call(*d, nch, drop, seed) for d, (_, _, _, nch, drop, seed) in zip(data, cases)]
    outs = [(torch.empty(B, H, device="cuda", dtype=torch.bfloat16), torch.empty(B, H, device="cuda", dtype=torch.bfloat16),
             torch.empty(B, H, device="cuda")) for (B, _, H, _, _, _) in cases]
    torch.cuda.synchronize()
    stats0 = torch.cuda.memory_stats()
    free0 = torch.cuda.mem_get_info()[0]
    for d, (_, _, _, nch, drop, seed), o in zip(data, cases, outs):       # the two calls, back to back on one stream
        _static(*d, nch, drop, seed, o)
    torch.cuda.synchronize()
    stats1 = torch.cuda.memory_stats()
    # nothing was allocated between the two synchronisations: not through torch's allocator, not from the driver
    assert stats1["allocation.all.allocated"] == stats0["allocation.all.allocated"]
    assert stats1["segment.all.allocated"] == stats0["segment.all.allocated"]
    assert torch.cuda.mem_get_info()[0] >= free0
    for (pooled, hrf, keep), (wp, wh, wk), (B, S, H, nch, drop, _) in zip(outs, want, cases):
        assert torch.isfinite(hrf.float()).all()
        assert torch.equal(pooled, wp) and torch.equal(hrf, wh) and torch.equal(keep, wk), \
            "B=%d S=%d H=%d nch=%d: max |diff| hrf %g" % (B, S, H, nch, float((hrf.float() - wh.float()).abs().max()))
        if drop > 0:
            dropped = float((keep == 0).float().mean())
            assert abs(dropped - drop) < 0.02


def test_static_scratch_pool_rejects_sizes_beyond_the_limit():
    from phantom_vlb_b200 import lib as _l
    B, S, H, nch = 8, 256, 4096, 17                                        # 8 * 17 * 4096 > 524 288
    d = _inputs(B, S, H, 7)
    hrf = torch.full((B, H), 3.0, device="cuda", dtype=torch.bfloat16)
    out = (torch.empty_like(hrf), hrf, torch.empty(B, H, device="cuda"))
    with pytest.raises(RuntimeError, match="exceeds the built-in scratch"):
        _static(*d, nch, 0.0, 1, out)
    torch.cuda.synchronize()
    assert bool((hrf == 3.0).all())                                        # nothing was launched
