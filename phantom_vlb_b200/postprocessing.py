"""Accuracy brain maps from validation metrics -- mirror of src/postprocessing/make_acc_brainmaps.py:33-58.

The reference reads Lightning's `metrics.csv`, keeps the validation rows (`val/brain_loss` not NaN), takes the columns
whose name contains `ROI` in sorted order, squares the per-parcel r and paints it into the atlas volume with nilearn's
`NiftiLabelsMasker.inverse_transform` (vol[x,y,z] = r2[label[x,y,z]-1], 0 outside the atlas), then saves NIfTI / an HTML
viewer. Here the gather runs on the GPU (`vlb_paint_labels`), the volume is written as `.npy` and, optionally, as an
uncompressed NIfTI-1 file whose 348-byte header is written by hand (nibabel / nilearn are not needed). The HTML viewer is
out of scope (browser-side JS).

`MetricsCSVWriter` is the columnar replacement for the P `self.log` calls of LogValAccuracyCallback (src/utils.py:108-110):
one row per validation epoch with the same `val_corr_ROI_%06d` / `val_corr_avg` column names, so the reference's
make_acc_brainmaps.py parses it unchanged.
"""
import csv
import os
import struct

import numpy as np
import torch

from . import ops


class MetricsCSVWriter:
    """Append-only `metrics.csv` in the layout of Lightning's CSVLogger: a header with the union of all keys seen so far,
    empty cells for metrics a row does not have (pandas reads them as NaN -- what `val_df[...].notna()` relies on)."""

    def __init__(self, log_dir, name="metrics.csv"):
        os.makedirs(log_dir, exist_ok=True)
        self.path = os.path.join(log_dir, name)
        self.columns, self.rows = [], []

    def append(self, row):
        for k in row:
            if k not in self.columns:
                self.columns.append(k)
        self.rows.append({k: (float(v) if not isinstance(v, (int, str)) else v) for k, v in row.items()})
        with open(self.path, "w", newline="") as f:
            w = csv.DictWriter(f, fieldnames=self.columns, restval="")
            w.writeheader()
            w.writerows(self.rows)

    def log_validation(self, epoch, step, val_loss, correlations):
        """correlations: 1-D array of per-parcel r (ONE device->host copy upstream, not P scalar syncs)."""
        c = np.asarray(correlations, dtype=np.float64)
        row = {"epoch": int(epoch), "step": int(step), "val/brain_loss": float(val_loss)}
        row.update({"val_corr_ROI_%06d" % i: c[i] for i in range(c.shape[0])})
        row["val_corr_avg"] = float(np.mean(c))
        self.append(row)


def read_validation_rois(metrics_path):
    """[n_val_epochs, P] per-parcel r from `<metrics_path>/metrics.csv` (make_acc_brainmaps.py:41-48)."""
    import pandas as pd
    df = pd.read_csv(os.path.join(metrics_path, "metrics.csv"), sep=",")
    val = df[df["val/brain_loss"].notna()]
    cols = sorted(c for c in val.columns if "ROI" in c)
    return val[cols].to_numpy(dtype=np.float64)


def write_nifti1(path, vol, affine=None):
    """Minimal single-file NIfTI-1 (.nii, float32, uncompressed): 348-byte header + 4 pad bytes + data in x-fastest order."""
    vol = np.asarray(vol, dtype=np.float32)
    assert vol.ndim == 3
    aff = np.eye(4, dtype=np.float32) if affine is None else np.asarray(affine, dtype=np.float32)
    hdr = bytearray(348)
    struct.pack_into("<i", hdr, 0, 348)                                     # sizeof_hdr
    struct.pack_into("<8h", hdr, 40, 3, vol.shape[0], vol.shape[1], vol.shape[2], 1, 1, 1, 1)   # dim
    struct.pack_into("<h", hdr, 70, 16)                                     # datatype FLOAT32
    struct.pack_into("<h", hdr, 72, 32)                                     # bitpix
    vox = [float(np.linalg.norm(aff[:3, i])) for i in range(3)]
    struct.pack_into("<8f", hdr, 76, 1.0, vox[0], vox[1], vox[2], 0.0, 0.0, 0.0, 0.0)           # pixdim
    struct.pack_into("<f", hdr, 108, 352.0)                                 # vox_offset
    struct.pack_into("<f", hdr, 112, 1.0)                                   # scl_slope
    struct.pack_into("<h", hdr, 254, 2)                                     # sform_code = aligned
    for r in range(3):
        struct.pack_into("<4f", hdr, 280 + 16 * r, *[float(v) for v in aff[r]])                   # srow_x/y/z
    hdr[344:348] = b"n+1\0"
    with open(path, "wb") as f:
        f.write(bytes(hdr))
        f.write(b"\0\0\0\0")
        f.write(np.asfortranarray(vol).tobytes(order="F"))


def make_brainmaps(metrics_path, atlas_labels, out_path, export_nii=False, affine=None, device="cuda"):
    """One r^2 volume per validation epoch. atlas_labels: integer volume, 0 = background, l in 1..P = parcel l-1
    (the label convention NiftiLabelsMasker uses for a labels image with consecutive labels). Returns the volumes."""
    rois = read_validation_rois(metrics_path)
    labels = torch.as_tensor(np.ascontiguousarray(atlas_labels), dtype=torch.int32, device=device)
    vols = []
    for i in range(rois.shape[0]):
        r = torch.as_tensor(rois[i], dtype=torch.float32, device=device)
        vol = ops.paint_labels(r, labels, square=True).cpu().numpy()
        np.save("%s_val-%d.npy" % (out_path, i), vol)
        if export_nii:
            write_nifti1("%s_val-%d.nii" % (out_path, i), vol, affine)
        vols.append(vol)
    return vols
