"""Ingest benchmark at BASELINE configs[1] sizes (16 x 768 x 1152 fp32 files, batch 8): GPU dataset
statistics, the synchronous batch path and the prefetching loader, with the reference's host path
(np.load + fp32 standardise per sample, one process) timed beside them.  Files live in a scratch
directory (page cache after the first pass); GB/s = payload bytes of the .npy files per second."""
import os, sys, tempfile, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
from bias_gan_b200 import stats as S
from bias_gan_b200.gpsro_dataset import GPSRODataset, PrefetchLoader

C, H, W, N, B = 16, 768, 1152, int(os.environ.get("NFILES", "32")), 8
dev = torch.device("cuda", 0)
with tempfile.TemporaryDirectory() as d:
    rng = np.random.default_rng(0)
    mu, sig = (np.arange(C) - 8.0).reshape(C, 1, 1).astype(np.float32), (1.0 + np.arange(C) / 4.0).reshape(C, 1, 1).astype(np.float32)
    for i in range(N):
        a = rng.standard_normal((C, H, W), dtype=np.float32) * sig + mu
        np.save(os.path.join(d, "data_in_%05d.npy" % i), a)
        np.save(os.path.join(d, "data_out_%05d.npy" % i), a + np.float32(0.1))
    payload = 2.0 * N * C * H * W * 4
    print("%d file pairs of %dx%dx%d fp32 = %.2f GB payload" % (N, C, H, W, payload / 1e9))
    for rep in range(2):
        t0 = time.perf_counter()
        st = S.summarize_dataset(d, os.path.join(d, "stats.npz"))
        torch.cuda.synchronize()
        dt = time.perf_counter() - t0
        print("stats.npz on the GPU (read + reduce), pass %d: %.3f s  %.2f GB/s" % (rep, dt, payload / dt / 1e9))
    t0 = time.perf_counter()
    tok = None
    for i in range(4):   # the reference script on the host: 4 files, scaled
        a = np.load(os.path.join(d, "data_in_%05d.npy" % i)).astype(np.float64)
        t = (1, np.average(a, axis=(1, 2)), np.average(np.square(a), axis=(1, 2)), np.amin(a, axis=(1, 2)), np.amax(a, axis=(1, 2)))
    dt = (time.perf_counter() - t0) / 4
    print("reference create_token on the host (numpy fp64, 1 core): %.3f s per file  %.2f GB/s" % (dt, C * H * W * 4 / dt / 1e9))

    ds = GPSRODataset(d, os.path.join(d, "stats.npz"), list(range(C)), normalization_type="MeanVariance",
                      num_intra_threads=int(os.environ.get("INTRA", "4")), read_device=dev, send_device=dev)
    nb = len(ds) // B
    for rep in range(2):
        torch.cuda.synchronize(); t0 = time.perf_counter()
        for i in range(nb):
            ds.get_batch_fields(list(range(i * B, (i + 1) * B)))
        torch.cuda.synchronize(); dt = time.perf_counter() - t0
        print("synchronous batches (read -> device -> stage), pass %d: %.1f ms/batch  %.2f GB/s" % (rep, dt / nb * 1e3, payload / dt / 1e9))
    loader = PrefetchLoader(ds, B, depth=2)
    for step_ms in (0, 130):
        cycles = int(step_ms * 1e-3 * 1.9e9)
        base = 0.0
        if step_ms:                                    # the same GPU work alone
            torch.cuda.synchronize(); t0 = time.perf_counter()
            for i in range(nb):
                torch.cuda._sleep(cycles)
            torch.cuda.synchronize(); base = (time.perf_counter() - t0) / nb
        torch.cuda.synchronize(); t0 = time.perf_counter(); blocked = 0.0
        it = iter(loader)
        for i in range(nb):
            tw = time.perf_counter()
            x, y, names = next(it)                     # host time blocked waiting for the producer
            blocked += time.perf_counter() - tw
            if step_ms:
                torch.cuda._sleep(cycles)
        torch.cuda.synchronize(); dt = time.perf_counter() - t0
        for _ in it:
            pass
        print("prefetching loader + %5.1f ms of GPU work per batch: %.1f ms/batch  %.2f GB/s  (host blocked in next(): %.1f ms/batch)"
              % (base * 1e3, dt / nb * 1e3, payload / dt / 1e9, blocked / nb * 1e3))
    # the reference's host path for one sample: np.load + H2D + fp32 standardise on the device
    cpu = GPSRODataset(d, os.path.join(d, "stats.npz"), list(range(C)), normalization_type="MeanVariance",
                       num_intra_threads=1, read_device=torch.device("cpu"), send_device=dev)
    torch.cuda.synchronize(); t0 = time.perf_counter()
    for i in range(B):
        cpu[i]
    torch.cuda.synchronize(); dt = time.perf_counter() - t0
    print("reference-style __getitem__ (host read, H2D, fp32 standardise), %d samples: %.1f ms/batch-of-8  %.2f GB/s"
          % (B, dt * 1e3, 2.0 * B * C * H * W * 4 / dt / 1e9))
