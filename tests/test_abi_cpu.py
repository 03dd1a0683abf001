"""C-ABI checks that need no GPU: the shared library loads and exports every symbol include/vlb_b200.h declares."""
import ctypes
import os

from phantom_vlb_b200 import lib


def test_header_parses_and_library_exports_every_symbol():
    protos = lib.parse_header()
    assert len(protos) >= 35
    for required in ("vlb_gemm_bf16", "vlb_attn_fwd", "vlb_attn_bwd", "vlb_hrf_pool_fwd", "vlb_pearson_update",
                     "vlb_adamw_step", "vlb_splice_embed", "vlb_weight_mask", "vlb_last_error_string"):
        assert required in protos
    if not os.path.exists(lib.LIB_PATH):
        from phantom_vlb_b200.build import build_lib
        build_lib(verbose=False)
    L = ctypes.CDLL(lib.LIB_PATH)
    for name in protos:
        assert hasattr(L, name), "symbol %s declared in the header but not exported" % name
    assert lib.lib().vlb_abi_version() == 1


def test_argument_errors_are_reported_without_a_gpu():
    # argument validation happens before any CUDA call: bad head_dim must come back as rc != 0 with a message
    rc = lib.lib().vlb_attn_fwd(None, None, None, 8, None, 8, None, 1, 128, 3, 2, 128, 1, 1.0, None)
    assert rc != 0 and b"attn_fwd" in lib.lib().vlb_last_error_string()


def test_no_fallback_when_library_missing(monkeypatch):
    import phantom_vlb_b200.lib as L
    monkeypatch.setattr(L, "_lib", None)
    monkeypatch.setattr(L, "LIB_PATH", "/nonexistent/libvlb_b200.so")
    try:
        L.lib()
    except RuntimeError as e:
        assert "no fallback" in str(e).lower()
    else:
        raise AssertionError("loading a missing library must fail loudly")
