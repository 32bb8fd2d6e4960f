"""`CamDataset` / `CamDataloader` of the reference's single-file CAM loader
(src/deepCam/data/cam_numpy_singlefile_dataset.py:15-74,131-213): one `.npy` of all samples in NHWC
`[N,768,1152,16]`, one of labels `[N,768,1152,2]`, read with `split_axis=True` readers.

Same constructors, attributes and yielded tuples.  With `preprocess=True` the reference yields the
strided view `data[..., channels].permute(0,3,1,2)` and `label[..., 0]`; so does this class.  The
B200 path adds `fields=True`: the batch is channel-selected, cast and laid out as the networks' bf16
space-to-depth field by one stage kernel straight from the reader's NHWC device tensor (the CAM
loader does not standardise: shift 0, scale 1), instead of materialising the permuted fp32 tensor.
`CamSoftwareDataset` (np.memmap on the host, :77-128) is the reference's conventional-IO baseline
and is not re-implemented here.
"""
import numpy as np
import torch
from torch.utils.data import Dataset

from . import kernels as K
from . import numpy_reader as nr


def _open(filename, device, inter=None, intra=None, batchsize=None):
    r = nr.numpy_reader(split_axis=True, device=device)
    if inter is not None:
        r.num_inter_threads = inter
    if intra is not None:
        r.num_intra_threads = intra
    r.parse(filename)
    if batchsize is not None:
        r.set_batchsize(batchsize)
    r.init_file(filename)
    return r


class CamDataset(Dataset):

    def __init__(self, filename_filenames, filename_data, filename_label, channels, device=-1, preprocess=True):
        self.channels = channels
        self.preprocess = preprocess
        self.files = list(np.load(filename_filenames))
        self.npr_data = _open(filename_data, device)
        self.npr_label = _open(filename_label, device)
        self.length = self.npr_data.num_samples
        assert self.length == self.npr_label.num_samples
        assert self.length == len(self.files)
        print("Initialized dataset with ", self.length, " samples.")

    def __del__(self):
        for r in (getattr(self, "npr_data", None), getattr(self, "npr_label", None)):
            if r is not None:
                r.finalize_file()

    def __len__(self):
        return self.length

    @property
    def shapes(self):
        return self.npr_data.shape, self.npr_label.shape

    def __getitem__(self, idx):
        X = self.npr_data.get_sample(idx)
        if self.preprocess:
            X = X[..., self.channels].permute(2, 0, 1)
        Y = self.npr_label.get_sample(idx)
        if self.preprocess:
            Y = torch.squeeze(Y[..., 0])
        return X, Y, self.files[idx]


class CamDataloader(object):

    def __init__(self, filename_filenames, filename_data, filename_label, channels, batchsize, shuffle=False,
                 num_inter_threads=1, num_intra_threads=1, device=-1, preprocess=True, fields=False):
        self.channels = channels
        self.preprocess = preprocess
        self.fields = fields
        self.device = device
        self.files = np.load(filename_filenames)
        self.shuffle = shuffle
        self.batchsize = batchsize
        self.npr_data = _open(filename_data, device, num_inter_threads, num_intra_threads, batchsize)
        self.npr_label = _open(filename_label, device, num_inter_threads, num_intra_threads, batchsize)
        self.length = self.npr_data.num_samples
        assert self.length == self.npr_label.num_samples
        assert self.length == self.files.shape[0]
        self.indices = range(0, self.length)
        if self.shuffle:
            self.indices = np.random.permutation(self.indices)     # the global numpy RNG, as in the reference
        if fields:
            if device < 0:
                raise RuntimeError("CamDataloader(fields=True) needs the reader's device path")
            C = len(channels)
            dev = torch.device("cuda", device)
            self._shift, self._scale = torch.zeros(C, device=dev), torch.ones(C, device=dev)
        print("Initialized data loader with ", self.length, " samples.")

    @property
    def shapes(self):
        return self.npr_data.shape, self.npr_label.shape

    def __del__(self):
        for r in (getattr(self, "npr_data", None), getattr(self, "npr_label", None)):
            if r is not None:
                r.finalize_file()

    def __len__(self):
        return self.length

    def __iter__(self):
        for index in range(0, self.length, self.batchsize):
            if (index + self.batchsize) > self.length:
                break
            batch = self.indices[index: index + self.batchsize]
            data = self.npr_data.get_batch(list(batch))
            if self.batchsize == 1:
                data = data.unsqueeze(0)
            if self.fields:      # one kernel: channel select + bf16 + field layout, no fp32 NCHW copy
                data = K.stage_normalize(data, "NHWC", list(self.channels), self._shift, self._scale, s2d=True)
            elif self.preprocess:
                data = data[..., self.channels].permute(0, 3, 1, 2)
            elif self.batchsize == 1:
                data = data.squeeze(0)                              # reference: no unsqueeze without preprocess
            label = self.npr_label.get_batch(list(batch))
            if self.preprocess or self.fields:
                label = torch.squeeze(label[..., 0])
                if self.batchsize == 1:
                    label = label.unsqueeze(0)
            filenames = self.files[batch]
            yield data, label, filenames
