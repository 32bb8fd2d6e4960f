"""Dataset statistics producer: the `stats.npz` the standardisation constants are read from
(reference: src/utils/summarize_data_gpsro.py — `create_token` :31-50, `merge_token` :5-27, the
driver loop :68-89; consumed by data/gpsro_dataset.py:87-116).

A token is `(n, mean[C], sqmean[C], min[C], max[C])` in fp64, exactly the reference's tuple.  Here the
per-file reduction runs on the GPU over the tensor the reader delivered (bgr_channel_stats, one read
of the payload); merging tokens is a few hundred bytes of fp64 arithmetic and stays on the host, in
the reference's order and with its formulas, so that a dataset summarised here and one summarised by
the reference script differ only by the summation order inside one file (<= 1e-12 relative).
"""
import os

import numpy as np
import torch

from . import kernels as K
from . import numpy_reader as nr


def create_tokens(batch, data_format="nchw"):
    """Tokens of every sample of a device batch [B,C,H,W] (nchw) / [B,H,W,C] (nhwc): a list of
    reference tuples (n = 1 each)."""
    out = K.channel_stats(batch, "NCHW" if data_format == "nchw" else "NHWC").cpu().numpy()
    return [(1, o[0].copy(), o[1].copy(), o[2].copy(), o[3].copy()) for o in out]


def create_token(sample, data_format="nchw"):
    """summarize_data_gpsro.py:31-50 for one sample already on the device ([C,H,W] or [H,W,C])."""
    return create_tokens(sample.unsqueeze(0).contiguous(), data_format)[0]


def merge_token(token1, token2):
    """summarize_data_gpsro.py:5-27: count-weighted means, element-wise min / max."""
    n1, mean1, sq1, lo1, hi1 = token1
    n2, mean2, sq2, lo2, hi2 = token2
    n = n1 + n2
    return (n, (n1 / n) * mean1 + (n2 / n) * mean2, (n1 / n) * sq1 + (n2 / n) * sq2,
            np.minimum(lo1, lo2), np.maximum(hi1, hi2))


def summarize_files(files, device=0, data_format="nchw", intra_threads=4):
    """The reference's loop (:68-82) over one file list: token of the first file, then
    `merge_token(create_token(file_i), running)` for i = 1.. in sorted order."""
    reader = nr.numpy_reader(False, device)
    reader.num_intra_threads = intra_threads
    running = None
    for f in files:
        reader.parse(f)
        reader.set_batchsize(1)
        reader.init_file(f)
        sample = reader.get_sample(0)
        reader.finalize_file()
        tok = create_token(sample, data_format)
        running = tok if running is None else merge_token(tok, running)
    return running


def summarize_dataset(root, out_path=None, device=0, data_format="nchw"):
    """stats.npz of a GPSRO-style directory (`data_in_*.npy` / `data_out_*.npy`), same keys as the
    reference (:86-89).  Returns the dict that was written."""
    names = sorted(os.listdir(root))
    data_files = [os.path.join(root, x) for x in names if x.endswith(".npy") and x.startswith("data_in_")]
    label_files = [os.path.join(root, x) for x in names if x.endswith(".npy") and x.startswith("data_out_")]
    if not data_files or len(data_files) != len(label_files):
        raise RuntimeError("summarize_dataset: need matching data_in_*.npy / data_out_*.npy files in " + root)
    d, l = summarize_files(data_files, device, data_format), summarize_files(label_files, device, data_format)
    assert d[0] == l[0]
    stats = dict(count=d[0], data_mean=d[1], data_sqmean=d[2], data_minval=d[3], data_maxval=d[4],
                 label_mean=l[1], label_sqmean=l[2], label_minval=l[3], label_maxval=l[4])
    if out_path is not None:
        np.savez(out_path, **stats)
    return stats
