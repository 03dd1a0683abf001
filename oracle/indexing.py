"""ORACLE (test infrastructure only).  Integer TR / frame / token indexing contract of the reference,
restated in plain numpy; every function cites the reference lines it follows.  Must be bit-exact."""
import math

import numpy as np


def frame_sample_uniform(duration, num_frames):
    """VideoLLaMA2/videollama2/mm_utils.py:106-121 (mode='uniform')."""
    seg_size = float(duration - 1) / num_frames
    ids = [(seg_size * i + seg_size * (i + 1)) / 2 for i in range(num_frames)]
    return np.round(np.array(ids) + 1e-6).astype(int)


def tr_end_times(duration, tr=1.49):
    """src/preprocessing/videollama2_vlb_extractfeatures.py:315."""
    return (np.array(range(1, math.ceil(duration / tr))) * tr).tolist()


def video_chunk_indices(end_time, win_dur, fps, num_frames_of_video, tr=1.49, frames_per_tr=4):
    """src/preprocessing/videollama2_vlb_extractfeatures.py:320-343 -> (f_start, f_end, num_frames, sampled indices,
    number of black pad frames)."""
    start_time = max(0, end_time - tr * win_dur)
    f_start = max(int(start_time * fps) - 1, 0)
    f_end = min(int(end_time * fps) - 1, num_frames_of_video - 1)
    all_idx = list(range(f_start, f_end + 1))
    num_frames = round((end_time - start_time) / tr) * frames_per_tr
    sampled = [all_idx[i] for i in frame_sample_uniform(len(all_idx), num_frames)]
    return f_start, f_end, num_frames, sampled, win_dur * frames_per_tr - len(sampled)


def text_budget(num_frames, max_len=2048, tok_per_frame=169):
    """src/preprocessing/videollama2_vlb_extractfeatures.py:198-212: L_text = max_len - T'*169 + 1, T' = T//2 + 1."""
    return max_len - (num_frames // 2 + 1) * tok_per_frame + 1


def visual_token_lags(num_frames, delay=3, window=3, tr=1.49):
    """src/preprocessing/videollama2_vlb_lazyloading.py:99,108-115: lag (s) of each of the T' visual token groups to
    the target TR mid-time."""
    tp = num_frames // 2 + 1
    step = window / (tp - 1)
    abs_tr_delay = (window - 1) + delay + 0.5
    return tr * (abs_tr_delay - np.arange(0, window + step, step))[:tp]


def tr_mid_time(row, delay=3, window=3, tr=1.49):
    """src/preprocessing/videollama2_vlb_lazyloading.py:99: time stamp (s) of aligned row `row`'s target TR, assigned to the
    middle of the BOLD slice."""
    return ((window - 1) + delay + 0.5 + row) * tr


def bold_alignment(n_feat_rows, n_bold_rows, delay=3, window=3):
    """src/preprocessing/videollama2_vlb_lazyloading.py:97-130: features[(window-1):], bold[(window-1)+delay:],
    n_rows = min of both -> (first feature row, first bold row, n_rows)."""
    f0, b0 = window - 1, (window - 1) + delay
    return f0, b0, min(n_feat_rows - f0, n_bold_rows - b0)


def weight_mask_row(pad_len, inst_len, dialog_len, vis_w, lang_w, tok_per_frame, max_len):
    """src/litmodule/videollama2_vlb_litmodule.py:178-203 for one sample, in float64 (no dtype rounding)."""
    trial = np.concatenate([np.repeat(np.asarray(vis_w, dtype=np.float64), tok_per_frame), np.zeros(2 + inst_len),
                            np.asarray(lang_w, dtype=np.float64)[:dialog_len], np.zeros(4 + pad_len)])
    return np.concatenate([np.zeros(max_len - trial.shape[0]), trial])
