"""`GPSRODataset` with the reference's constructor, attributes and `__getitem__`
(src/deepCam/data/gpsro_dataset.py:17-175), plus what the B200 path adds on top of it: whole batches
read straight into one device buffer and standardised + cast + laid out as bf16 fields by one stage
kernel per tensor (`get_batch_fields`), and a prefetching loader that keeps the next batches in flight
on a second stream (`PrefetchLoader`) in place of
`DataLoader(train_set, local_batch_size, drop_last=True)` (gpsro_train/train_gan.py:199).

File order, sharding and the shift/scale vectors are the reference's to the bit (init_files :19-33,
statistics :87-116).  Deliberate differences: `masks=False` means no masks (the reference tests
`self.masks is not None`, which is true for its own default `False`, and then needs `masks_*` files);
the field path has no mixing augmentation (`augmentation_mode` must be None there).
"""
import os
import threading

import numpy as np
import torch
from torch.utils.data import Dataset

from . import kernels as K
from . import numpy_reader as nr


class GPSRODataset(Dataset):

    def init_files(self, source):
        self.source = source
        self.allfiles = sorted([x.replace("data_in_", "") for x in os.listdir(self.source)
                                if x.endswith(".npy") and x.startswith("data_in_")])
        if self.shuffle:
            self.rng.shuffle(self.allfiles)
        shard_size = len(self.allfiles) // self.shard_num   # remainder files are dropped, as in the reference
        start = shard_size * self.shard_idx
        self.files = self.allfiles[start:start + shard_size]
        self.length = len(self.files)

    def _reader(self, prefix, devindex, num_intra_threads):
        r = nr.numpy_reader(split_axis=False, device=devindex)
        r.num_intra_threads = num_intra_threads
        r.parse(os.path.join(self.source, prefix + self.files[0]))
        return r

    def __init__(self, source, statsfile, channels, normalization_type="MinMax", shuffle=True, masks=False,
                 augmentation_mode=None, shard_idx=0, shard_num=1, num_intra_threads=1, seed=12345,
                 read_device=torch.device("cpu"), send_device=torch.device("cpu")):
        self.channels = channels
        self.normalization_type = normalization_type
        self.shuffle = shuffle
        self.masks = masks
        self.seed = seed
        self.rng = np.random.RandomState(seed)
        self.shard_idx = shard_idx
        self.shard_num = shard_num
        self.read_device = torch.device(read_device)
        self.send_device = torch.device(send_device)
        self.augmentation_mode = augmentation_mode
        self.init_files(source)
        if self.length == 0:
            raise RuntimeError("GPSRODataset: no data_in_*.npy files for this shard in " + str(source))
        if self.augmentation_mode == "static":
            self.mixing = self.rng.rand(self.length)
        devindex = -1 if self.read_device.type == "cpu" else (self.read_device.index or 0)
        self.npr_data = self._reader("data_in_", devindex, num_intra_threads)
        self.npr_label = self._reader("data_out_", devindex, num_intra_threads)
        if self.masks:
            self.npr_mask = self._reader("masks_", devindex, num_intra_threads)
        stats = np.load(statsfile)
        ch = self.channels
        if self.normalization_type == "MinMax":
            data_shift = stats["data_minval"][ch]
            data_scale = 1. / (stats["data_maxval"][ch] - data_shift)
            label_shift = stats["label_minval"][ch]
            label_scale = 1. / (stats["label_maxval"][ch] - label_shift)
        elif self.normalization_type == "MeanVariance":
            data_shift = stats["data_mean"][ch]
            data_scale = 1. / np.sqrt(stats["data_sqmean"][ch] - np.square(data_shift))
            label_shift = stats["label_mean"][ch]
            label_scale = 1. / np.sqrt(stats["label_sqmean"][ch] - np.square(label_shift))
        else:
            raise NotImplementedError("normalization_type " + str(self.normalization_type))
        as_t = lambda v: torch.tensor(np.reshape(v, (v.shape[0], 1, 1)).astype(np.float32),
                                      requires_grad=False).to(self.send_device)
        self.data_shift, self.data_scale = as_t(data_shift), as_t(data_scale)
        self.label_shift, self.label_scale = as_t(label_shift), as_t(label_scale)
        print("Initialized dataset with ", self.length, " samples.")

    def __len__(self):
        return self.length

    @property
    def shapes(self):
        return self.npr_data.shape, self.npr_label.shape

    def _read_one(self, reader, prefix, idx, out=None):
        path = os.path.realpath(os.path.join(self.source, prefix + self.files[idx]))
        reader.init_file(path)
        t = reader.get_sample(0, out=out)
        reader.finalize_file()
        return t

    def __getitem__(self, idx):
        """The reference's per-sample contract: fp32 [C_file,H,W] tensors, standardised with the selected
        channels' constants (the reference broadcasts [len(channels),1,1] against the file's channels)."""
        data = self._read_one(self.npr_data, "data_in_", idx)
        label = self._read_one(self.npr_label, "data_out_", idx)
        mask = self._read_one(self.npr_mask, "masks_", idx) if self.masks else None
        if data.device != self.send_device:
            data = data.to(self.send_device)
        if label.device != self.send_device:
            label = label.to(self.send_device)
        if self.masks and mask.device != self.send_device:
            mask = mask.to(self.send_device)
        if self.augmentation_mode == "static":
            data = data + self.mixing[idx] * label
            label = (1. - self.mixing[idx]) * label
        elif self.augmentation_mode == "dynamic":
            p = self.rng.rand()
            data = data + p * label
            label = (1. - p) * label
        data = self.data_scale * (data - self.data_shift)
        label = self.label_scale * (label - self.label_shift)
        if self.masks:
            return data, label, mask, self.files[idx]
        return data, label, self.files[idx]

    # ---- the B200 path: whole batches, one stage kernel per tensor ------------------------------
    def _raw_buffers(self, batch):
        shp_d, shp_l = self.shapes
        if self.read_device.type != "cuda":
            raise RuntimeError("get_batch_fields needs read_device = the GPU (the reader's device path)")
        mk = lambda shp: torch.empty([batch] + list(shp), dtype=torch.float32, device=self.read_device)
        return mk(shp_d), mk(shp_l)

    def get_batch_fields(self, indices, raw=None, out=None, s2d=True):
        """(data, label, filenames) for the samples `indices`: bf16 fields
        [B, H/2, W/2, 4*len(channels)] (space-to-depth, the networks' input layout) or NHWC with
        s2d=False.  Every file is read straight into its slot of one [B,C,H,W] device buffer and the
        batch is standardised, channel-selected, cast and laid out by one kernel per tensor
        (replaces __getitem__ x B + default collate)."""
        if self.augmentation_mode is not None:
            raise NotImplementedError("mixing augmentation is not available on the field path")
        B = len(indices)
        raw = raw if raw is not None else self._raw_buffers(B)
        raw_d, raw_l = raw[0], raw[1]
        for b, idx in enumerate(indices):
            self._read_one(self.npr_data, "data_in_", idx, out=raw_d[b])
            self._read_one(self.npr_label, "data_out_", idx, out=raw_l[b])
        out_d, out_l = (out[0], out[1]) if out is not None else (None, None)
        ch = list(self.channels)
        x = K.stage_normalize(raw_d, "NCHW", ch, self.data_shift.flatten(), self.data_scale.flatten(), s2d=s2d, out=out_d)
        y = K.stage_normalize(raw_l, "NCHW", ch, self.label_shift.flatten(), self.label_scale.flatten(), s2d=s2d, out=out_l)
        if self.masks:   # the reference yields (data, label, masks, filename) then (gpsro_dataset.py:172-175)
            raw_m = raw[2] if len(raw) > 2 else torch.empty_like(raw_d)
            for b, idx in enumerate(indices):
                self._read_one(self.npr_mask, "masks_", idx, out=raw_m[b])
            zero = torch.zeros(len(ch), dtype=torch.float32, device=raw_m.device)
            m = K.stage_normalize(raw_m, "NCHW", ch, zero, torch.ones_like(zero), s2d=s2d,
                                  out=out[2] if (out is not None and len(out) > 2) else None)   # masks are not standardised
            return x, y, m, [self.files[i] for i in indices]
        return x, y, [self.files[i] for i in indices]


class PrefetchLoader:
    """Batches of a GPSRODataset in index order (`drop_last=True`, like the reference's DataLoader),
    produced `depth` batches ahead by a worker thread on its own CUDA stream: file reads, host->device
    copies and the stage kernels of batch i+1.. overlap the training step of batch i.  Each yielded
    batch lives in one of `depth + 1` preallocated slots; a slot is rewritten only after the
    consumer's stream has passed the point where it asked for the batch after next, so a yielded
    batch stays valid until the next `__next__` returns."""

    def __init__(self, dataset, batch_size, depth=2, s2d=True):
        if dataset.read_device.type != "cuda" or dataset.send_device != dataset.read_device:
            raise RuntimeError("PrefetchLoader needs read_device == send_device == the GPU")
        self.ds, self.B, self.depth, self.s2d = dataset, int(batch_size), int(depth), s2d
        self.dev = dataset.read_device
        self.nbatches = len(dataset) // self.B
        self.stream = torch.cuda.Stream(device=self.dev)
        nslots = self.depth + 1
        shp = dataset.shapes[0]
        C, H, W = len(dataset.channels), shp[1], shp[2]
        fshape = (self.B, H // 2, W // 2, 4 * C) if s2d else (self.B, H, W, C)
        ntens = 3 if dataset.masks else 2     # data, label (, masks): the reference's batch tuple
        self.raw = [dataset._raw_buffers(self.B) for _ in range(nslots)]
        if dataset.masks:
            self.raw = [r + (torch.empty_like(r[0]),) for r in self.raw]
        self.fields = [tuple(torch.empty(fshape, dtype=torch.bfloat16, device=self.dev) for _ in range(ntens))
                       for _ in range(nslots)]
        self.ready = [torch.cuda.Event() for _ in range(nslots)]      # producer -> consumer
        self.released = [None] * nslots                               # consumer -> producer

    def __len__(self):
        return self.nbatches

    def _produce(self, queue_put):
        try:
            with torch.cuda.stream(self.stream):
                for i in range(self.nbatches):
                    slot = i % len(self.fields)
                    self.free[slot].acquire()                 # the consumer has handed the slot back
                    if self.released[slot] is not None:
                        self.stream.wait_event(self.released[slot])
                    idx = list(range(i * self.B, (i + 1) * self.B))
                    names = self.ds.get_batch_fields(idx, raw=self.raw[slot], out=self.fields[slot], s2d=self.s2d)[-1]
                    self.ready[slot].record(self.stream)
                    queue_put((slot, names, None))
            queue_put((None, None, None))
        except BaseException as e:                            # surfaced in the consumer
            queue_put((None, None, e))

    def __iter__(self):
        import queue
        q = queue.Queue()
        self.free = [threading.Semaphore(1) for _ in self.fields]
        worker = threading.Thread(target=self._produce, args=(q.put,), daemon=True)
        worker.start()
        held = []                                             # slots the consumer may still be using
        while True:
            slot, names, err = q.get()
            if err is not None:
                raise err
            if slot is None:
                break
            cur = torch.cuda.current_stream(self.dev)
            cur.wait_event(self.ready[slot])
            held.append(slot)
            if len(held) > 1:                                 # the batch before the previous one is done with
                old = held.pop(0)
                ev = torch.cuda.Event()
                ev.record(cur)
                self.released[old] = ev
                self.free[old].release()
            yield self.fields[slot] + (names,)
        worker.join()
