"""ORACLE (test infrastructure only).  nilearn's Glover-HRF regressor, restated step by step.

The reference computes every HRF weight with (src/utils.py:30-37)
    compute_regressor(exp_condition=[[0],[1],[1]], hrf_model="glover", frame_times=[0.0, time_diff])[0][-1, 0]
nilearn is a third-party dependency that is NOT under /root/reference and NOT installed here:
    pinned version     : nilearn==0.12.0 (requirements_rorqual.txt:41)
    call sites         : src/utils.py:10,30-35; lazyloading.py:113-115,135-137
    restated algorithm : nilearn/glm/first_level/hemodynamic_models.py -- compute_regressor -> _sample_condition (event
                         boxcar on an oversampled grid that starts min_onset=-24 s before the first frame), glover_hrf ->
                         _gamma_difference_hrf (gamma pdfs: delay 6, undershoot 12, dispersion 0.9 / 0.9, ratio 0.35, loc=dt,
                         32 s long, normalised to unit SUM), np.convolve truncated to the regressor length,
                         _resample_regressor (linear interpolation at the frame times), _orthogonalize (no-op for 1 column)
PARITY UNPINNED: nilearn cannot be executed here, so this file follows the published algorithm; the weights it produces
are only INPUTS of the training step (vis_weights / lang_weights of the batch contract).
"""
import numpy as np
from scipy.interpolate import interp1d
from scipy.stats import gamma

OVERSAMPLING = 50
MIN_ONSET = -24.0
TIME_LENGTH = 32.0


def glover_hrf(t_r, oversampling=OVERSAMPLING, time_length=TIME_LENGTH, onset=0.0):
    dt = t_r / oversampling
    time_stamps = np.linspace(0, time_length, np.rint(float(time_length) / dt).astype(int))
    time_stamps -= onset
    delay, undershoot, dispersion, u_dispersion, ratio = 6.0, 12.0, 0.9, 0.9, 0.35
    peak_gamma = gamma.pdf(time_stamps, delay / dispersion, loc=dt, scale=dispersion)
    undershoot_gamma = gamma.pdf(time_stamps, undershoot / u_dispersion, loc=dt, scale=u_dispersion)
    hrf = peak_gamma - ratio * undershoot_gamma
    hrf /= hrf.sum()
    return hrf


def sample_condition(exp_condition, frame_times, oversampling=OVERSAMPLING, min_onset=MIN_ONSET):
    n = frame_times.size
    n_hr = ((n - 1) * 1.0 / (frame_times.max() - frame_times.min())
            * (frame_times.max() * (1 + 1.0 / (n - 1)) - frame_times.min() - min_onset) * oversampling) + 1
    hr_frame_times = np.linspace(frame_times.min() + min_onset, frame_times.max() * (1 + 1.0 / (n - 1)),
                                 np.rint(n_hr).astype(int))
    onsets, durations, values = tuple(map(np.asanyarray, exp_condition))
    tmax = len(hr_frame_times)
    regressor = np.zeros_like(hr_frame_times).astype(np.float64)
    t_onset = np.minimum(np.searchsorted(hr_frame_times, onsets), tmax - 1)
    for t, v in zip(t_onset, values):
        regressor[t] += v
    t_offset = np.minimum(np.searchsorted(hr_frame_times, onsets + durations), tmax - 1)
    for i, t in enumerate(t_offset):
        if t < (tmax - 1) and t == t_onset[i]:
            t_offset[i] += 1
    for t, v in zip(t_offset, values):
        regressor[t] -= v
    return np.cumsum(regressor), hr_frame_times


def compute_regressor_glover(exp_condition, frame_times):
    frame_times = np.asarray(frame_times, dtype=np.float64)
    t_r = float(frame_times.max()) / (np.size(frame_times) - 1)
    hr_regressor, hr_frame_times = sample_condition(exp_condition, frame_times)
    h = glover_hrf(t_r)
    conv_reg = np.array([np.convolve(hr_regressor, h)[: hr_regressor.size]])
    f = interp1d(hr_frame_times, conv_reg)
    return f(frame_times).T


def get_hrf_weight(time_diff):
    """src/utils.py:14-37."""
    reg = compute_regressor_glover(np.array([[0], [1], [1]]), np.array([0.0, time_diff]))
    return float(reg[-1, 0])
