#!/usr/bin/env python
"""bench.py — device-timed G+D training samples/s of the Bias-GAN hot path on synthetic
16x768x1152 fields (BASELINE.json configs[1]: 1xB200, bf16 GAN train step, batch 8/GPU).

    python bench.py --gpus 1 --steps 5 --warmup 3
    python -m torch.distributed.run --nproc-per-node N ... bench.py --gpus N ...
    python bench.py --config hires ...      # BASELINE.json configs[4]: 16x1536x2304 fields, batch 4/GPU
    python bench.py --config ingest ...     # BASELINE.json configs[3]: .npy files -> reader -> stage (-> step), reader GB/s
    python bench.py --impl reference ...    # the reference's CPU path (its own modules where /root/reference
                                            # exists, else the oracle port) at the stated config

One step = stage (standardise -> bf16 field) of inputs and targets + one G+D step
(train_gan.py:250-298: D phase + G phase, AdamW).  `value` times it with the raw fp32 batch
resident in HBM; `e2e` adds the pinned-host -> device copy of the raw batch and the device ->
host read of both losses every step.  Prints ONE JSON line on rank 0.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

C, H, W = 16, 768, 1152
F_G, F_D = 627.85e9, 380.30e9                  # forward FLOPs per sample at 768x1152
W_MIN = 4 * F_G + 8 * F_D                      # useful step work per sample = 5553.8 GFLOP
METRIC = "GAN train samples/sec (G+D step)"
# BASELINE.json configs: [1] = baseline, [4] = hires (4x the pixels: FLOPs per sample scale exactly x4), [3] = ingest
CONFIGS = {"baseline": (768, 1152, 8), "hires": (1536, 2304, 4), "ingest": (768, 1152, 8)}


def workload_string(config, batch):
    """The same string in the product arm and the reference arm: it names BASELINE.json's configuration."""
    return ("%s config: G+D train step (train_gan.py:250-298), 16x%dx%d fields, local batch %d, "
            "DeepLabv3+/Xception G (Interpolate) + Xception D, AdamW" % (config, H, W, batch))


def set_config(name, batch=None):
    """Select the workload; returns the local batch."""
    global H, W, W_MIN
    H, W, b = CONFIGS[name]
    W_MIN = (4 * F_G + 8 * F_D) * (H * W) / (768.0 * 1152.0)
    return batch or b


def peaks():
    try:
        return json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json"))), "measured"
    except Exception:
        return {"hbm_gbs": 6650.0, "bf16_tflops": 1590.0, "bf16_tflops_sustained": 1400.0}, "fallback"


class ClockSampler(threading.Thread):
    """nvidia-smi clocks / throttle reasons sampled while the timed region runs."""
    Q = "clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown," \
        "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index, self.rows, self._halt = index, [], threading.Event()

    def run(self):
        while not self._halt.is_set():
            try:
                out = subprocess.run(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q,
                                      "--format=csv,noheader,nounits"], capture_output=True, text=True, timeout=5).stdout
                f = [v.strip() for v in out.strip().split(",")]
                if len(f) >= 6:
                    self.rows.append(f)
            except Exception:
                pass
            self._halt.wait(0.2)

    def stop(self):
        self._halt.set()
        self.join(timeout=6)
        if not self.rows:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        sm = sorted(int(r[0]) for r in self.rows if r[0].isdigit())
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = [n for i, n in enumerate(names) if any(r[2 + i].lower().startswith("active") for r in self.rows)]
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": int(self.rows[0][1]) if self.rows[0][1].isdigit() else None,
                "reasons": reasons, "samples": len(self.rows)}


REF_SRC = "/root/reference/src/deepCam"


def _reference_modules(h, w):
    """(G, D, GANLoss, step_fn, kind): the reference's OWN modules where /root/reference exists (the build
    container; conv2d_local stubbed and D's linear resized as tests/golden/make_model_golden.py does, SURVEY
    F1/F4), else the oracle port of them (the GPU box has no /root/reference).  Checker code: only the
    cpu_baseline leg and --impl reference come here."""
    import torch
    import torch.nn as nn
    from oracle import ref_model as R
    if os.path.isdir(REF_SRC):
        try:
            import types
            sys.modules.setdefault("conv2d_local", types.ModuleType("conv2d_local"))
            if REF_SRC not in sys.path:
                sys.path.insert(0, REF_SRC)
            from architecture.gpsro import deeplab_gan as dxg
            from utils.losses import GANLoss
            G = dxg.Generator(C, C, "Interpolate", "Normal", 0, os=16, normalizer=nn.BatchNorm2d)
            D = dxg.Discriminator(C, os=16, normalizer=nn.BatchNorm2d)
            D.linear = nn.Linear(2048 * -(-h // 16) * -(-w // 16), 1)
            return G, D, GANLoss, "reference"
        except Exception as e:   # fall through to the port
            sys.stderr.write("bench: reference modules not importable (%r); using the oracle port\n" % (e,))
    return R.GeneratorRef(C, C), R.DiscriminatorRef(C, (h, w)), R.GANLossRef, "port"


def cpu_step_sample(steps, warmup, threads, h, w, b=2):
    """One G+D step (train_gan.py:250-298, reference schedule) of the reference CPU path: fp32, all host
    threads, local batch 2 (BatchNorm needs > 1) on an h x w field.  Returns (samples/s scaled to
    full fields of the selected config, seconds per step, kind, description)."""
    import torch
    from oracle import ref_model as R
    from oracle.seeded import seeded_fields
    torch.set_num_threads(threads)
    torch.manual_seed(333)
    G, D, GANLoss, kind = _reference_modules(h, w)
    G.train(); D.train()
    go = torch.optim.AdamW(G.parameters(), lr=1e-4, eps=1e-8, weight_decay=1e-5)
    do = torch.optim.AdamW(D.parameters(), lr=1e-4, eps=1e-8, weight_decay=1e-5)
    crit = GANLoss("ModifiedMinMax", b, torch.device("cpu"))
    x, y = seeded_fields(1, b, C, h, w)
    for _ in range(warmup):
        R.gan_step_ref(G, D, go, do, crit, x, y.clone())
    t0 = time.perf_counter()
    for _ in range(steps):
        R.gan_step_ref(G, D, go, do, crit, x, y.clone())
    dt = (time.perf_counter() - t0) / steps
    frac = (h * w) / float(H * W)
    what = "full %dx%d field" % (h, w) if frac == 1.0 else "%dx%d crop (1/%d field), scaled to full-field samples/s" % (h, w, round(1 / frac))
    desc = "%s, fp32 G+D step (reference schedule), local batch %d, %s" % (
        "the reference's own modules" if kind == "reference" else "oracle port of the reference modules", b, what)
    return b * frac / dt, dt, kind, desc


def _write_gpsro_files(d, n, with_stats=True):
    """n synthetic file pairs data_in_/data_out_ (N(0,1)*sigma_c + mu_c, label = data + 0.1) and
    stats.npz by the reference's token algorithm (closed form here: the files share one noise plane)."""
    import numpy as np
    rng = np.random.default_rng(1000)
    mu = (np.arange(C) - 8.0).reshape(C, 1, 1).astype(np.float32)
    sig = (1.0 + np.arange(C) / 4.0).reshape(C, 1, 1).astype(np.float32)
    base = rng.standard_normal((C, H, W), dtype=np.float32) * sig + mu
    for i in range(n):
        a = base + np.float32(0.001 * i)
        np.save(os.path.join(d, "data_in_%05d.npy" % i), a)
        np.save(os.path.join(d, "data_out_%05d.npy" % i), a + np.float32(0.1))
    if with_stats:
        z = base.astype(np.float64)
        m, q = z.mean(axis=(1, 2)), (z * z).mean(axis=(1, 2))
        np.savez(os.path.join(d, "stats.npz"), data_mean=m, data_sqmean=q, data_minval=z.min(axis=(1, 2)),
                 data_maxval=z.max(axis=(1, 2)), label_mean=m + 0.1, label_sqmean=q + 0.2 * m + 0.01,
                 label_minval=z.min(axis=(1, 2)) + 0.1, label_maxval=z.max(axis=(1, 2)) + 0.1)
    return 2.0 * n * C * H * W * 4   # payload bytes


def cpu_reader_legs(d, nfiles, threads):
    """SURVEY §8(d) CPU legs (i) and (ii), on the files in `d` (page cache):
    (i) the reference numpy_reader's CPU path (oracle/_ref/numpy_reader.so, device=-1, intra threads 1 and 8;
        absent -> the numpy restatement oracle/npy_oracle.py, kind "port") + the reference normalisation
        scale*(x - shift) in torch CPU fp32 (gpsro_dataset.py:169-170);
    (ii) the reference's np.memmap host path (cam_numpy_singlefile_dataset.py:96-125): memmap at offset 128,
        np.copy, NHWC -> NCHW transpose."""
    import numpy as np
    import torch
    torch.set_num_threads(threads)
    out = {"cores": threads}
    files = [os.path.join(d, "data_in_%05d.npy" % i) for i in range(min(nfiles, 4))]
    nbytes = float(C * H * W * 4)
    shift = torch.arange(C, dtype=torch.float32).view(C, 1, 1) - 8.0
    scale = 1.0 / (1.0 + torch.arange(C, dtype=torch.float32).view(C, 1, 1) / 4.0)
    ref = None
    so = os.path.join(ROOT, "oracle", "_ref", "numpy_reader.so")
    if os.path.exists(so):
        try:
            import importlib.util
            spec = importlib.util.spec_from_file_location("numpy_reader", so)
            ref = importlib.util.module_from_spec(spec)
            spec.loader.exec_module(ref)
        except Exception as e:
            sys.stderr.write("bench: oracle/_ref/numpy_reader.so did not load (%r)\n" % (e,))
            ref = None
    legs = {}
    if ref is not None:
        for intra in (1, 8):
            r = ref.numpy_reader(split_axis=False, device=-1)
            r.num_intra_threads = intra
            r.parse(files[0])
            t_read = t_norm = 0.0
            for f in files:
                t0 = time.perf_counter()
                r.init_file(f); x = r.get_sample(0); r.finalize_file()
                t1 = time.perf_counter()
                x = scale * (x - shift)
                t_read += t1 - t0
                t_norm += time.perf_counter() - t1
            legs["intra%d" % intra] = {"read_gbs": len(files) * nbytes / t_read / 1e9,
                                       "read_plus_normalise_gbs": len(files) * nbytes / (t_read + t_norm) / 1e9,
                                       "ms_per_sample": (t_read + t_norm) / len(files) * 1e3}
        out["reader"] = {"kind": "reference", "what": "reference numpy_reader (oracle/_ref), device=-1, + scale*(x-shift) in torch CPU fp32", **legs}
    else:
        from oracle import npy_oracle as orc
        t_read = t_norm = 0.0
        for f in files:
            t0 = time.perf_counter()
            x = torch.from_numpy(orc.read_samples(f, [0], False))
            t1 = time.perf_counter()
            x = scale * (x.view(C, H, W) - shift)
            t_read += t1 - t0
            t_norm += time.perf_counter() - t1
        out["reader"] = {"kind": "port", "what": "numpy restatement of the reference reader (oracle/npy_oracle.py) + scale*(x-shift) in torch CPU fp32",
                         "read_gbs": len(files) * nbytes / t_read / 1e9,
                         "read_plus_normalise_gbs": len(files) * nbytes / (t_read + t_norm) / 1e9,
                         "ms_per_sample": (t_read + t_norm) / len(files) * 1e3}
    # (ii) the CAM single-file host path: one NHWC file of 2 samples
    nhwc = os.path.join(d, "cam_data.npy")
    a = np.load(files[0])
    np.save(nhwc, np.ascontiguousarray(np.stack([a, a]).transpose(0, 2, 3, 1)))
    t0 = time.perf_counter()
    mm = np.memmap(nhwc, dtype=np.float32, mode="r", offset=128, shape=(2, H, W, C))
    for i in range(2):
        x = np.transpose(np.copy(mm[i]), (2, 0, 1))
        x = np.ascontiguousarray(x)
    dt = (time.perf_counter() - t0) / 2
    out["memmap"] = {"kind": "reference-style", "what": "np.memmap(offset=128) + np.copy + NHWC->NCHW transpose, one core "
                     "(cam_numpy_singlefile_dataset.py:96-125)", "gbs": nbytes / dt / 1e9, "ms_per_sample": dt * 1e3}
    os.remove(nhwc)
    return out


def run_reference(args):
    """--impl reference: the reference's CPU implementation of the path on the host cores.  One G+D step at the
    stated configuration's field size (full 16x768x1152, local batch 2 -- BatchNorm rejects 1 -- reported per
    sample) is the measurement; it takes about a minute on 16 cores, so it is timed ONCE whatever --steps says
    (a crop-scaled step, which flatters the CPU with cache-resident activations, is the labelled fallback when
    BGR_REF_SAMPLE=crop or the host has too little memory)."""
    if int(os.environ.get("RANK", "0")) != 0:
        return
    b = set_config(args.config if args.config != "ingest" else "baseline", args.batch)
    threads = os.cpu_count() or 1
    full = os.environ.get("BGR_REF_SAMPLE", "full") != "crop"
    try:
        import psutil
        if psutil.virtual_memory().available < 48 * 2 ** 30:
            full = False
    except Exception:
        pass
    if full:
        v, dt, kind, sample = cpu_step_sample(1, 0, threads, H, W)
        sample += "; one step timed once (about a minute of CPU work), --steps/--warmup not repeated"
    else:
        v, dt, kind, sample = cpu_step_sample(max(1, args.steps), min(args.warmup, 1), threads, H // 4, W // 4)
    line = {"impl": "reference", "metric": METRIC, "value": v, "unit": "samples/s", "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": dt * 1e3 * b / 2.0, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": workload_string(args.config if args.config != "ingest" else "baseline", b),
                       "global_batch": b * max(1, args.gpus), "parallelism": "dp%d" % max(1, args.gpus),
                       "reference_arm": "the reference's CPU implementation of the same step on the host cores; BatchNorm "
                                        "rejects batch 1 and a batch-%d step needs %d x the memory, so one step at local "
                                        "batch 2 is timed and reported per sample" % (b, b // 2)},
            "cpu_baseline": {"value": v, "unit": "samples/s", "cores": threads, "kind": kind, "sample": sample,
                             "seconds_per_step_of_2": dt},
            "e2e": {"value": v, "unit": "samples/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    emit(line)


def run_torch_gpu(args, autocast):
    """Same-GPU comparator ("the bar to beat", not the CPU baseline and not the product):
    the oracle port of the reference modules through stock PyTorch/cuDNN on cuda:0, the reference's own
    step schedule (train_gan.py:250-298), fp32 or bf16 autocast, device-timed like the product arm."""
    import torch
    from oracle import ref_model as R
    dev = torch.device("cuda", 0)
    torch.backends.cudnn.benchmark = True
    g = torch.Generator().manual_seed(1)
    line = {"impl": "torch-bf16-autocast" if autocast else "torch-fp32", "metric": METRIC, "unit": "samples/s", "n_gpus": 1,
            "steps": args.steps, "warmup": args.warmup, "higher_is_better": True, "data": "synthetic",
            "dtype": "bf16 autocast" if autocast else "f32 (TF32 off)"}
    torch.backends.cudnn.allow_tf32 = False
    torch.backends.cuda.matmul.allow_tf32 = False
    for b in (args.batch, args.batch // 2, args.batch // 4):
        G = D = go = do = None
        try:
            torch.manual_seed(333)
            G, D = R.GeneratorRef(C, C).to(dev).train(), R.DiscriminatorRef(C, (H, W)).to(dev).train()
            opt = lambda m: torch.optim.AdamW(m.parameters(), lr=1e-4, eps=1e-8, weight_decay=1e-5, fused=True)
            go, do = opt(G), opt(D)
            crit = R.GANLossRef("ModifiedMinMax", b, dev)
            x, y = torch.randn(b, C, H, W, generator=g).to(dev), torch.randn(b, C, H, W, generator=g).to(dev)

            def step():
                with torch.autocast("cuda", dtype=torch.bfloat16, enabled=autocast):
                    R.gan_step_ref(G, D, go, do, crit, x, y)
            for _ in range(max(args.warmup, 2)):
                step()
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(args.steps):
                step()
            e1.record()
            torch.cuda.synchronize()
            ms = e0.elapsed_time(e1) / args.steps
            line.update({"value": b / (ms * 1e-3), "ms_per_step": ms, "peak_mem_gib": torch.cuda.max_memory_allocated() / 2 ** 30,
                         "config": {"workload": "reference modules (oracle port) via PyTorch/cuDNN on 1xB200, reference step "
                                                "schedule, 16x768x1152 fields, local batch %d" % b}})
            break
        except torch.cuda.OutOfMemoryError:
            import gc
            G = D = go = do = step = None
            gc.collect()
            torch.cuda.empty_cache()
            line["note"] = "batch %d does not fit in HBM with stock PyTorch" % b
    emit(line)


def drop_page_cache():
    """Best effort (root on the GPU box): make the next read of the files come from storage, not RAM."""
    try:
        os.sync()
        with open("/proc/sys/vm/drop_caches", "w") as f:
            f.write("3\n")
        return True
    except Exception:
        return False


def ingest_measure(torch, dev, d, B, step_fn=None, fields_out=None, intra=4, passes=2, cold=False):
    """The ingest half of the path on the files in `d`: (a) reader only -- every file pair through
    bgr_reader_read (pread -> pinned ring -> per-chunk cudaMemcpyAsync) into one [B,C,H,W] device buffer;
    (b) prefetching loader = reader + stage kernels, two batches ahead on its own stream; (c) loader feeding
    `step_fn(x_field, y_field)` (the graphed G+D step).  GB/s = payload bytes of the .npy files (headers
    excluded) delivered to device memory per second of wall time."""
    from bias_gan_b200.gpsro_dataset import GPSRODataset, PrefetchLoader
    ds = GPSRODataset(d, os.path.join(d, "stats.npz"), list(range(C)), normalization_type="MeanVariance",
                      num_intra_threads=intra, read_device=dev, send_device=dev)
    nb = len(ds) // B
    payload = 2.0 * nb * B * C * H * W * 4
    res = {"files": 2 * nb * B, "payload_gb": payload / 1e9, "intra_threads": intra,
           "path": "bgr_reader_read: pread -> pinned ring -> cudaMemcpyAsync per chunk -> device (no cuFile: the box has no nvidia-fs)"}
    raw_d, raw_l = ds._raw_buffers(B)

    def reader_pass():
        torch.cuda.synchronize(dev)
        t0 = time.perf_counter()
        for i in range(nb):
            for b in range(B):
                ds._read_one(ds.npr_data, "data_in_", i * B + b, out=raw_d[b])
                ds._read_one(ds.npr_label, "data_out_", i * B + b, out=raw_l[b])
        torch.cuda.synchronize(dev)
        return time.perf_counter() - t0

    if cold and drop_page_cache():
        res["reader_gbs_cold"] = payload / reader_pass() / 1e9
    else:
        reader_pass()
    res["reader_gbs"] = payload / min(reader_pass() for _ in range(passes)) / 1e9
    loader = PrefetchLoader(ds, B, depth=2)

    def loader_pass(fn):
        torch.cuda.synchronize(dev)
        t0 = time.perf_counter()
        out = None
        for x, y, _names in loader:
            if fn is not None:
                out = fn(x, y)
        if out is not None:
            float(out[0])   # read a loss back: the host waits for the last step
        torch.cuda.synchronize(dev)
        return time.perf_counter() - t0

    res["loader_gbs"] = payload / min(loader_pass(None) for _ in range(passes)) / 1e9
    if step_fn is not None:
        dt = min(loader_pass(step_fn) for _ in range(passes))
        res["with_step"] = {"samples_per_s": nb * B / dt, "ms_per_batch": dt / nb * 1e3, "reader_gbs": payload / dt / 1e9,
                            "what": "files -> PrefetchLoader (reader + stage) -> graphed G+D step, %d batches" % nb}
    del loader, ds
    return res


def host_path_measure(torch, dev, d, n=4):
    """The reference-style host path for the same files (gpsro_dataset.py:130-170 semantics): read on the
    host, .to(device), fp32 standardise on the device, one sample at a time."""
    from bias_gan_b200.gpsro_dataset import GPSRODataset
    cpu = GPSRODataset(d, os.path.join(d, "stats.npz"), list(range(C)), normalization_type="MeanVariance",
                       num_intra_threads=1, read_device=torch.device("cpu"), send_device=dev)
    cpu[0]
    torch.cuda.synchronize(dev)
    t0 = time.perf_counter()
    for i in range(n):
        cpu[i]
    torch.cuda.synchronize(dev)
    dt = time.perf_counter() - t0
    return {"gbs": 2.0 * n * C * H * W * 4 / dt / 1e9, "ms_per_sample": dt / n * 1e3,
            "what": "per-sample host read + H2D + fp32 scale*(x-shift) on the device (reference __getitem__ semantics)"}


def conv_roofline_step(step_fn, torch, K):
    """Dominant kernel family (tcgen05 implicit GEMM): one kernel-by-kernel step records the signature
    of every conv launch (shapes, pixel strides, taps, epilogue options); each distinct signature is
    then timed alone with CUDA events on the launching stream, L2 flushed before every launch, and
    weighted by its launch count.  (Event pairs around the launches of a live step would time the
    host: launching ~3400 kernels one by one is slower than the GPU runs them.)"""
    sigs = {}
    orig_f, orig_w = K.conv_fprop, K.conv_wgrad

    def f_wrap(x, w, dy, dx, out=None, bias=None, want_stats=False):
        key = ("fprop", tuple(x.shape), x.stride(2), w.shape[0], tuple(dy), tuple(dx), bias is not None, bool(want_stats))
        sigs[key] = sigs.get(key, 0) + 1
        return orig_f(x, w, dy, dx, out=out, bias=bias, want_stats=want_stats)

    def w_wrap(x, gy, dy, dx, dw=None):
        key = ("wgrad", tuple(x.shape), x.stride(2), gy.shape[-1], tuple(dy), tuple(dx), False, False)
        sigs[key] = sigs.get(key, 0) + 1
        return orig_w(x, gy, dy, dx, dw=dw)

    K.conv_fprop, K.conv_wgrad = f_wrap, w_wrap
    try:
        step_fn()
        torch.cuda.synchronize()
    finally:
        K.conv_fprop, K.conv_wgrad = orig_f, orig_w
    dev = torch.device("cuda", torch.cuda.current_device())
    flush = torch.empty(256 * 1024 * 1024, dtype=torch.uint8, device=dev)
    by, reps, table = {}, 3, []
    for (kind, shape, ldx, cout, dy, dx, has_bias, has_stats), count in sigs.items():
        n, h, w_, cin = shape
        x = torch.randn(n, h, w_, max(ldx, cin), device=dev).bfloat16()[..., :cin]
        nt = len(dy)
        if kind == "fprop":
            wt = torch.randn(cout, nt, K.pad_channels(cin) if cin > 64 else cin, device=dev).bfloat16()[..., :cin]
            bias = torch.zeros(cout, device=dev) if has_bias else None
            y = K.act_empty((n, h, w_, cout), dev)
            fn = lambda: orig_f(x, wt, list(dy), list(dx), out=y, bias=bias, want_stats=has_stats)
        else:
            gy = K.act_empty((n, h, w_, cout), dev)
            gy.normal_()
            dwt = torch.zeros(cout, nt, cin, device=dev)
            fn = lambda: orig_w(x, gy, list(dy), list(dx), dw=dwt)
        fn()
        ms = 0.0
        for _ in range(reps):
            flush.zero_()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(); fn(); e1.record()
            torch.cuda.synchronize()
            ms += e0.elapsed_time(e1) / reps
        px_ = n * h * w_
        table.append((count * ms, count, ms * 1e3, kind, "%dx%dx%d" % (n, h, w_), cin, cout, nt, max(abs(v) for v in dy + dx),
                      "stats" if has_stats else ("bias" if has_bias else "-"), 2.0 * px_ * cin * cout * nt / (ms * 1e-3) / 1e12,
                      (2.0 * px_ * (cin + cout) + cout * nt * cin * (2.0 if kind == "fprop" else 4.0)) / (ms * 1e-3) / 1e9))
        a = by.setdefault(kind, [0.0, 0.0, 0, 0.0])
        a[0] += count * 2.0 * n * h * w_ * cin * cout * nt
        a[1] += count * ms
        a[2] += count
        # algorithmic bytes: every operand and the result once (bf16 activations/weights, fp32 dW)
        px = n * h * w_
        a[3] += count * (2.0 * px * (cin + cout) + cout * nt * cin * (2.0 if kind == "fprop" else 4.0))
        del x, fn
    path = os.environ.get("BGR_CONV_TABLE")   # per-signature table of the step's implicit-GEMM launches (profiles/)
    if path:
        with open(path, "w") as f:
            f.write("# every distinct implicit-GEMM launch signature of one G+D step, timed alone (CUDA events, L2 flushed), by\n"
                    "# share of the step's conv time.  GB/s = algorithmic bytes (operands + result once) / time: the HBM-bound\n"
                    "# signatures are the ones near the measured copy bandwidth, the tensor-bound ones near the bf16 peak.\n")
            f.write("#  total_ms  count   us/launch  kind   pixels        Cin  Cout taps maxoff epilogue  TFLOP/s    GB/s\n")
            for r in sorted(table, reverse=True):
                f.write("%10.3f %6d %11.1f  %-5s  %-12s %5d %5d %4d %6d %-8s %8.1f %7.0f\n" % r)
    flops = sum(v[0] for v in by.values())
    ms = sum(v[1] for v in by.values())
    return flops, ms, sum(v[2] for v in by.values()), \
        {k: {"tflops": v[0] / v[1] / 1e9, "ms": v[1], "launches": v[2], "algorithmic_bytes": v[3]} for k, v in by.items()}


def conv_traffic(n_launches):
    """DRAM bytes per conv launch from the committed ncu capture of this same step (profiles/):
    measured once per round under ncu, never inside a timed run."""
    try:
        t = json.load(open(os.path.join(ROOT, "profiles", "r02_conv_traffic.json")))
    except (OSError, ValueError):
        return None, None
    if t.get("launches_per_step") != n_launches:   # the capture is of another schedule: do not quote it
        return None, "profiles/r02_conv_traffic.json is stale (%s launches per step there)" % t.get("launches_per_step")
    return t["dram_bytes_per_launch"], "profiles/r02_conv_traffic.json"


_OUT = None


def _claim_stdout():
    """The driver reads ONE JSON line from stdout; NCCL (its version banner) and other native libraries
    write to fd 1 directly.  Keep a private handle on the real stdout for the result line and point fd 1 at
    stderr for everything else."""
    global _OUT
    if _OUT is None:
        sys.stdout.flush()
        _OUT = os.fdopen(os.dup(1), "w")
        os.dup2(2, 1)


def emit(line):
    _claim_stdout()
    _OUT.write(json.dumps(line) + "\n")
    _OUT.flush()


def main():
    _claim_stdout()
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200")
    ap.add_argument("--batch", type=int, default=None, help="local batch per GPU (default: the config's)")
    ap.add_argument("--config", default="baseline", choices=sorted(CONFIGS),
                    help="baseline = BASELINE.json configs[1] (16x768x1152, batch 8/GPU); hires = configs[4] "
                         "(16x1536x2304, batch 4/GPU); ingest = configs[3] (files -> reader -> stage -> step, reader GB/s)")
    ap.add_argument("--files", type=int, default=None, help="file pairs per rank for the ingest legs")
    ap.add_argument("--cold", action="store_true", help="ingest: also time a first pass after dropping the page cache")
    ap.add_argument("--no-step", action="store_true", help="ingest: reader + stage only, no training step behind them")
    ap.add_argument("--no-ingest", action="store_true", help="skip the reader legs of the default line")
    ap.add_argument("--no-grad-sync", action="store_true",
                    help="control run for the scaling analysis: N ranks, gradient all-reduce disabled (NOT data-parallel "
                         "training; separates NCCL time from the slowest-GPU power-cap effect)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--faithful", action="store_true", help="reference-faithful schedule (no lean shortcuts)")
    ap.add_argument("--no-graph", action="store_true", help="launch the step kernel by kernel instead of replaying "
                                                            "the captured CUDA graph")
    ap.add_argument("--profile", action="store_true",
                    help="warm up, then run --steps steps between cudaProfilerStart/Stop and exit (for ncu "
                         "--profile-from-start off); prints no bench value")
    args = ap.parse_args()
    if args.impl in ("torch-bf16", "torch-fp32"):
        args.batch = set_config(args.config if args.config != "ingest" else "baseline", args.batch)
        return run_torch_gpu(args, args.impl == "torch-bf16")
    if args.impl == "reference":
        return run_reference(args)

    import torch
    from bias_gan_b200 import _lib, kernels as K
    from bias_gan_b200.comm import comm as Comm
    from bias_gan_b200.deeplab_gan import Discriminator, Generator
    from bias_gan_b200.losses import GANLoss, L1LossField
    from bias_gan_b200.step import GraphedGanStep, gan_step
    import torch.distributed as dist

    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device — the product path has no CPU fallback")
    cm = Comm()
    rank, world, lrank = cm.rank(), cm.size(), cm.local_rank()
    torch.cuda.set_device(lrank)
    dev = torch.device("cuda", lrank)
    B = set_config(args.config, args.batch)
    L = _lib.lib()
    L.bgr_launch_count.restype = __import__("ctypes").c_ulonglong

    # models / optimisers exactly as the run script (AdamW lr 1e-4, eps 1e-8, wd 1e-5), seed 333+7*rank
    torch.manual_seed(333 + 7 * rank)
    G = Generator(C, C, "Interpolate", "Normal", 0, os=16).to(dev).train()
    D = Discriminator(C, os=16, input_size=(H, W)).to(dev).train()
    if world > 1:
        cm.DistributedModel(G); cm.DistributedModel(D)
        if args.no_grad_sync:
            G.grad_sync.enabled = D.grad_sync.enabled = False
    g_opt = torch.optim.AdamW(G.parameters(), lr=1e-4, eps=1e-8, weight_decay=1e-5, fused=True, capturable=True)
    d_opt = torch.optim.AdamW(D.parameters(), lr=1e-4, eps=1e-8, weight_decay=1e-5, fused=True, capturable=True)
    crit_gan, crit_reg = GANLoss("ModifiedMinMax", B, dev), L1LossField()

    # synthetic raw batch: x = N(0,1)*sigma_c + mu_c, label = x + 0.1*noise; fp32 NCHW
    g = torch.Generator(device=dev).manual_seed(1000 + rank)
    mu = (torch.arange(C, device=dev, dtype=torch.float32) - 8.0)
    sigma = 1.0 + torch.arange(C, device=dev, dtype=torch.float32) / 4.0
    raw_x = torch.randn(B, C, H, W, device=dev, generator=g) * sigma.view(1, C, 1, 1) + mu.view(1, C, 1, 1)
    raw_y = raw_x + 0.1 * torch.randn(B, C, H, W, device=dev, generator=g)
    shift, scale = mu.contiguous(), (1.0 / sigma).contiguous()
    host_x, host_y = raw_x.cpu().pin_memory(), raw_y.cpu().pin_memory()
    dev_x, dev_y = torch.empty_like(raw_x), torch.empty_like(raw_y)

    graphed = None
    if not args.no_graph:
        # the whole G+D step (forward, backward, all-reduce hooks, AdamW) as one CUDA graph
        graphed = GraphedGanStep(G, D, g_opt, d_opt, crit_gan, crit_reg, (B, H // 2, W // 2, 4 * C), dev,
                                 lean=not args.faithful)

    def run_from(sx, sy):
        """stage (standardise -> bf16 field) + one G+D step from raw fp32 device tensors."""
        if graphed is not None:   # stage straight into the graph's static field buffers
            K.stage_normalize(sx, "NCHW", None, shift, scale, s2d=True, out=graphed.inputs)
            K.stage_normalize(sy, "NCHW", None, shift, scale, s2d=True, out=graphed.real)
            return graphed()
        xf = K.stage_normalize(sx, "NCHW", None, shift, scale, s2d=True)
        yf = K.stage_normalize(sy, "NCHW", None, shift, scale, s2d=True)
        return gan_step(G, D, g_opt, d_opt, crit_gan, crit_reg, xf, yf, lean=not args.faithful)

    def step(from_host=False):
        return run_from(raw_x, raw_y)

    def files_step(xf, yf):
        return graphed(xf, yf) if graphed is not None else gan_step(G, D, g_opt, d_opt, crit_gan, crit_reg, xf, yf,
                                                                      lean=not args.faithful)

    if args.config == "ingest":
        return run_ingest(args, torch, dist, cm, dev, B, None if args.no_step else files_step, step, L, graphed)

    # end-to-end: every step copies its raw batch from pinned host memory and reads both losses back.
    # The copy of step i+1 runs on a copy stream into the other of two device buffers while step i
    # computes (what a prefetching loader does); all K copies are inside the timed region.
    copy_stream = torch.cuda.Stream(device=dev)
    bufs = [(dev_x, dev_y), (torch.empty_like(raw_x), torch.empty_like(raw_y))]

    def issue_copy(k, after):
        with torch.cuda.stream(copy_stream):
            if after is not None:
                copy_stream.wait_event(after)
            bufs[k][0].copy_(host_x, non_blocking=True)
            bufs[k][1].copy_(host_y, non_blocking=True)
            ev = torch.cuda.Event()
            ev.record(copy_stream)
        return ev

    def e2e_steps(n):
        main = torch.cuda.current_stream()
        t0 = torch.cuda.Event()
        t0.record(main)
        ready, staged, out = issue_copy(0, t0), [None, None], None
        for i in range(n):
            k = i & 1
            main.wait_event(ready)
            if graphed is not None:
                K.stage_normalize(bufs[k][0], "NCHW", None, shift, scale, s2d=True, out=graphed.inputs)
                K.stage_normalize(bufs[k][1], "NCHW", None, shift, scale, s2d=True, out=graphed.real)
                xf = yf = None
            else:
                xf = K.stage_normalize(bufs[k][0], "NCHW", None, shift, scale, s2d=True)
                yf = K.stage_normalize(bufs[k][1], "NCHW", None, shift, scale, s2d=True)
            staged[k] = torch.cuda.Event()
            staged[k].record(main)
            if i + 1 < n:   # buffer k^1 was last read by the stage kernels of step i-1
                ready = issue_copy(k ^ 1, staged[k ^ 1])
            if graphed is not None:
                dl, gl = graphed()
            else:
                dl, gl = gan_step(G, D, g_opt, d_opt, crit_gan, crit_reg, xf, yf, lean=not args.faithful)
            out = (float(dl), float(gl))   # D2H read of both losses (host sync, as metric_average does)
        return out

    def eager_step():
        xf = K.stage_normalize(raw_x, "NCHW", None, shift, scale, s2d=True)
        yf = K.stage_normalize(raw_y, "NCHW", None, shift, scale, s2d=True)
        return gan_step(G, D, g_opt, d_opt, crit_gan, crit_reg, xf, yf, lean=not args.faithful)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(n, from_host):
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        c0 = L.bgr_launch_count()
        e0.record()
        if from_host:
            e2e_steps(n)
        else:
            for _ in range(n):
                step(False)
        e1.record()
        barrier()
        ms = torch.tensor([e0.elapsed_time(e1)], device=dev)
        if world > 1:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        launches = (L.bgr_launch_count() - c0) // max(n, 1)
        if graphed is not None:   # the stage kernels counted above + the kernels inside the replayed graph
            launches += graphed.launches_per_step
        return float(ms) / n, launches

    for _ in range(max(args.warmup, 3)):
        step(False)
    if args.profile:
        torch.cuda.synchronize()
        torch.cuda.profiler.start()
        for _ in range(args.steps):
            step(False)
        torch.cuda.synchronize()
        torch.cuda.profiler.stop()
        if rank == 0:
            emit({"profile_only": True, "steps": args.steps})
        return
    sampler = ClockSampler(lrank)
    if rank == 0:
        sampler.start()
    ms_step, launches = timed(args.steps, False)
    clocks = sampler.stop() if rank == 0 else None
    e2e_steps(2)
    ms_e2e, _ = timed(args.steps, True)
    peak_mem = torch.cuda.max_memory_allocated() / 2 ** 30

    if rank != 0:
        return
    pk, pk_src = peaks()
    value = world * B / (ms_step * 1e-3)
    e2e = world * B / (ms_e2e * 1e-3)
    for m in (G, D):   # rank 0 alone runs the instrumented step: no collectives in it
        if hasattr(m, "grad_sync"):
            m.grad_sync.enabled = False
    flops, ms_conv, n_conv, by_kind = conv_roofline_step(eager_step, torch, K)  # kernel by kernel, not the graph
    ach = flops / (ms_conv * 1e-3) / 1e12
    traffic, traffic_src = conv_traffic(n_conv) if (args.config == "baseline" and B == 8) else (None, "captured for the baseline config only")
    alg_bytes = sum(v["algorithmic_bytes"] for v in by_kind.values()) / max(n_conv, 1)
    line = {"metric": METRIC, "value": value, "unit": "samples/s", "n_gpus": world, "steps": args.steps,
            "warmup": max(args.warmup, 3), "ms_per_step": ms_step, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "bf16", "data": "synthetic",
            "config": {"workload": workload_string(args.config, B),
                       "global_batch": B * world, "parallelism": "dp%d" % world,
                       "schedule": "faithful" if args.faithful else "lean (result-identical, SURVEY F8)",
                       "grad_sync": "DISABLED (control run, not data-parallel training)" if args.no_grad_sync else
                                    ("bucketed NCCL all-reduce inside the graph" if world > 1 else "single rank"),
                       "launch": "eager" if graphed is None else "one CUDA graph replay per step",
                       "l2": "working set per step (>30 GB) far exceeds the 126 MB L2; no flush needed"},
            "clocks": clocks, "gpu_launches": int(launches),
            "e2e": {"value": e2e, "unit": "samples/s", "ms_per_step": ms_e2e,
                    "h2d_bytes_per_step": int(host_x.numel() * 4 * 2), "d2h_bytes_per_step": 8,
                    "pipeline": "pinned-host -> device copy of step i+1 on a copy stream while step i computes"},
            "roofline": {"bound": "tensor", "kernel": "conv_fprop_kernel/conv_wgrad_kernel (tcgen05 implicit GEMM)",
                         "achieved": ach, "peak": pk["bf16_tflops_sustained"], "unit": "TFLOP/s",
                         "frac": ach / pk["bf16_tflops_sustained"],
                         "frac_of_burst": ach / pk["bf16_tflops"], "peak_burst": pk["bf16_tflops"],
                         "regime": "every launch signature is timed alone (L2 flushed), which the recipe puts against the "
                                   "burst peak (frac_of_burst); inside the long step the kernels run at the sustained clock, "
                                   "which is what frac uses",
                         "traffic": traffic, "traffic_unit": "DRAM bytes per launch (mean over the step's conv launches)",
                         "traffic_source": traffic_src, "algorithmic_bytes_per_launch": alg_bytes,
                         "peak_source": pk_src + " (sustained)",
                         "launches_per_step": n_conv, "conv_ms_per_step": ms_conv, "by_kind": by_kind},
            "step_roofline": {"useful_gflop_per_sample": W_MIN / 1e9, "achieved_tflops": value / world * W_MIN / 1e12,
                              "frac_of_sustained_bf16": value / world * W_MIN / 1e12 / pk["bf16_tflops_sustained"]},
            "peak_mem_gib": peak_mem}
    threads = os.cpu_count() or 1
    if not args.no_ingest:
        # the ingest half on a driver-timed path: .npy files -> reader -> stage -> step (BASELINE "reader GB/s")
        import shutil
        import tempfile
        d = tempfile.mkdtemp(prefix="bgr_bench_")
        try:
            _write_gpsro_files(d, args.files or 4 * B)
            # rank 0 alone is here: with several ranks the captured step contains the gradient all-reduce, so the
            # step-fed leg (e2e.from_files) is measured at N = 1 only; `--config ingest` runs it on every rank
            rd = ingest_measure(torch, dev, d, B, step_fn=files_step if world == 1 else None)
            rd["host_path"] = host_path_measure(torch, dev, d)
            line["reader"] = rd
            line["reader_gbs"] = rd["reader_gbs"]
            line["e2e"]["from_files"] = rd.get("with_step")
            if not args.no_cpu_baseline:
                line.setdefault("cpu_baseline", {})["ingest"] = cpu_reader_legs(d, args.files or 4 * B, threads)
        finally:
            shutil.rmtree(d, ignore_errors=True)
    if not args.no_cpu_baseline:
        # bounded sample: one reference-schedule step at local batch 2 on a quarter-size field (10-30 s of CPU work)
        q = 2 if args.config == "baseline" else 4
        v, dt, kind, sample = cpu_step_sample(1, 0, threads, H // q, W // q)
        line.setdefault("cpu_baseline", {}).update({"value": v, "unit": "samples/s", "cores": threads, "kind": kind,
                                                    "sample": sample, "seconds": dt})
    emit(line)


def run_ingest(args, torch, dist, cm, dev, B, files_step, resident_step, L, graphed):
    """--config ingest (BASELINE.json configs[3]): per rank `--files` synthetic file pairs written to local
    scratch in setup (outside the timed region); timed = reader-only, loader (reader + stage) and loader ->
    graphed step.  One JSON line: value = aggregate reader GB/s over all ranks behind the training step
    (or of the loader alone with --no-step), max-over-ranks wall time."""
    import shutil
    import tempfile
    rank, world = cm.rank(), cm.size()
    nfiles = args.files or 4 * B
    d = tempfile.mkdtemp(prefix="bgr_ingest_r%d_" % rank)
    try:
        _write_gpsro_files(d, nfiles)
        for _ in range(max(args.warmup, 3)):
            resident_step(False)
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        sampler = ClockSampler(cm.local_rank())
        if rank == 0:
            sampler.start()
        rd = ingest_measure(torch, dev, d, B, step_fn=files_step, cold=args.cold, passes=max(1, min(args.steps, 3)))
        clocks = sampler.stop() if rank == 0 else None
        key = "with_step" if files_step is not None else None
        mine = rd["with_step"]["reader_gbs"] if key else rd["loader_gbs"]
        vals = torch.tensor([mine, rd["reader_gbs"], rd["loader_gbs"], rd.get("reader_gbs_cold", 0.0),
                             rd["with_step"]["samples_per_s"] if key else 0.0], device=dev, dtype=torch.float64)
        slow = vals.clone()
        if world > 1:   # aggregate = world x the slowest rank (every rank moves the same bytes)
            dist.all_reduce(slow, op=dist.ReduceOp.MIN)
        agg = (slow * world).tolist()
        if rank != 0:
            return
        host = host_path_measure(torch, dev, d)
        cpu = None if args.no_cpu_baseline else cpu_reader_legs(d, nfiles, os.cpu_count() or 1)
        line = {"metric": "reader GB/s (.npy payload delivered to device memory)", "value": agg[0], "unit": "GB/s",
                "n_gpus": world, "steps": args.steps, "warmup": max(args.warmup, 3), "higher_is_better": True,
                "scaling": "weak", "vs_baseline": None, "dtype": "f32 payload -> bf16 fields", "data": "synthetic",
                "ms_per_step": rd["with_step"]["ms_per_batch"] if key else None,
                "config": {"workload": "ingest config: %d file pairs of 16x%dx%d fp32 per rank (page cache%s) -> "
                                       "bgr_reader_read -> stage kernels%s, local batch %d"
                                       % (nfiles, H, W, " + one cold pass" if args.cold else "",
                                          " -> graphed G+D step" if key else "", B),
                           "parallelism": "dp%d (each rank reads its own files; no collective on the data path)" % world},
                "clocks": clocks,
                "gpu_launches": int(2 + (graphed.launches_per_step if (graphed is not None and key) else 0)),
                "reader_gbs": {"behind_step": agg[0] if key else None, "reader_only": agg[1], "loader": agg[2],
                               "reader_only_cold": agg[3] or None},
                "samples_per_s_from_files": agg[4] or None,
                "rank0": rd, "host_path": host,
                "e2e": {"value": agg[0], "unit": "GB/s", "h2d_bytes_per_step": int(2 * B * C * H * W * 4), "d2h_bytes_per_step": 4,
                        "note": "files -> pinned staging -> device inside the timed region; one loss read back per pass"},
                "roofline": {"bound": "hbm", "kernel": "stage_kernel", "unit": "GB/s", "achieved": None, "peak": peaks()[0]["hbm_gbs"],
                             "frac": None, "traffic": None,
                             "note": "the ingest path is bound by host reads + PCIe (about 50 GB/s per GPU), not by HBM; the "
                                     "stage kernel's own roofline is in the default config's line"},
                "cpu_baseline": cpu}
        emit(line)
    finally:
        shutil.rmtree(d, ignore_errors=True)


if __name__ == "__main__":
    main()
