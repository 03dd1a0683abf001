"""Host-side pieces of the brain-map path that need no GPU: the columnar metrics.csv writer (same column names as the
reference's per-ROI `self.log` calls, src/utils.py:108-110), the reader that mirrors make_acc_brainmaps.py:41-48, and the
hand-written NIfTI-1 header."""
import struct

import numpy as np

from phantom_vlb_b200.postprocessing import MetricsCSVWriter, read_validation_rois, write_nifti1


def test_metrics_csv_round_trip_matches_reference_parser(tmp_path):
    w = MetricsCSVWriter(str(tmp_path))
    rng = np.random.RandomState(0)
    w.append({"epoch": 0, "step": 10, "train/brain_loss": 1.25})          # a training row: no val columns
    r0, r1 = rng.rand(12) * 2 - 1, rng.rand(12) * 2 - 1
    w.log_validation(0, 20, 0.9, r0)
    w.append({"epoch": 1, "step": 30, "train/brain_loss": 1.0})
    w.log_validation(1, 40, 0.8, r1)
    rois = read_validation_rois(str(tmp_path))                            # val rows only, ROI columns in sorted order
    assert rois.shape == (2, 12)
    assert np.allclose(rois[0], r0) and np.allclose(rois[1], r1)
    import pandas as pd
    df = pd.read_csv(tmp_path / "metrics.csv")
    assert list(df.columns[:3]) == ["epoch", "step", "train/brain_loss"]
    assert "val_corr_ROI_000011" in df.columns and "val_corr_avg" in df.columns
    assert df["val/brain_loss"].notna().sum() == 2
    assert abs(df["val_corr_avg"].dropna().iloc[1] - r1.mean()) < 1e-12


def test_nifti1_header_and_payload(tmp_path):
    vol = np.arange(2 * 3 * 4, dtype=np.float32).reshape(2, 3, 4)
    aff = np.diag([2.0, 2.0, 2.5, 1.0]).astype(np.float32)
    aff[:3, 3] = [-90, -126, -72]
    p = tmp_path / "v.nii"
    write_nifti1(str(p), vol, aff)
    raw = p.read_bytes()
    assert len(raw) == 352 + vol.size * 4
    assert struct.unpack_from("<i", raw, 0)[0] == 348 and raw[344:348] == b"n+1\0"
    assert struct.unpack_from("<8h", raw, 40)[:4] == (3, 2, 3, 4)
    assert struct.unpack_from("<h", raw, 70)[0] == 16 and struct.unpack_from("<h", raw, 72)[0] == 32
    assert struct.unpack_from("<4f", raw, 76)[1:4] == (2.0, 2.0, 2.5)
    assert struct.unpack_from("<4f", raw, 280) == (2.0, 0.0, 0.0, -90.0)
    data = np.frombuffer(raw, dtype="<f4", offset=352).reshape(4, 3, 2).transpose(2, 1, 0)    # x fastest on disk
    assert np.array_equal(data, vol)


def test_vectorised_hrf_weights_match_scalar():
    from phantom_vlb_b200.utils import get_hrf_weight, hrf_weights
    lags = np.concatenate([1.49 * (5.5 - np.arange(0, 3 + 0.75, 0.75)), [-1.0, 0.0, 0.3, 30.0]])     # visual-token lags
    w = hrf_weights(lags)
    assert w.dtype == np.float64 and w.shape == lags.shape
    assert np.allclose(w, [get_hrf_weight(float(t)) for t in lags], rtol=1e-12, atol=0)
    assert w[-4] == 0.0 and w[-3] == 0.0 and 0.25 < hrf_weights([5.0])[0] < 0.3          # nilearn regressor: unit-sum HRF x 1 s event
