"""Data-parallel plumbing with the reference's `comm` surface (src/deepCam/comm/distributed.py:10-204):
rank/size/local_rank, metric_average, broadcast, printr, DistributedModel, DistributedOptimizer.

One process per GPU, torch.distributed over NCCL (gloo on CPU for tests).  DistributedModel does
not wrap the module (state-dict keys stay free of `module.` prefixes); it broadcasts parameters
from rank 0 like apex DDP does at construction and attaches a bucketed gradient all-reduce that
is launched from post-accumulate-grad hooks while backward is still running (the all-reduce runs
on NCCL's own stream and overlaps the remaining dgrad/wgrad kernels), averaged by world size.
"""
import os

import torch
import torch.distributed as dist


class BucketedAllReduce:
    """Gradient buckets of ~bucket_mb MB in reverse parameter order (≈ the order backward produces
    them, apex DDP's default `message_size` bucketing, comm/distributed.py:197).

    Protocol per optimizer update: `start()` -> loss.backward() (hooks launch a bucket's all-reduce as soon
    as its last gradient has been accumulated, on NCCL's stream, while backward continues) -> `finish()`
    -> optimizer.step().  `start()` drops whatever an earlier backward left behind, so a backward pass
    that touches this model's parameters without belonging to its update (the reference schedule
    back-propagates d_loss into G and g_loss into D) can never leak stale buckets into the
    next update; gan_step also disables the sync of the model that is not being stepped.

    One bucket = one flat fp32 tensor: gradients are gathered with one multi-tensor copy, averaged by
    the collective itself (ReduceOp.AVG on NCCL; sum and a pre-scale elsewhere) and scattered back with
    one multi-tensor copy -- no per-parameter kernels."""

    def __init__(self, module, bucket_mb=32, group=None):
        self.group = group
        self.world = dist.get_world_size(group) if dist.is_initialized() else 1
        params = [p for p in module.parameters()][::-1]
        cap = bucket_mb * 1024 * 1024 // 4
        self.buckets, cur, n = [], [], 0
        for p in params:
            cur.append(p)
            n += p.numel()
            if n >= cap:
                self.buckets.append(cur)
                cur, n = [], 0
        if cur:
            self.buckets.append(cur)
        self.bucket_of = {}
        for bi, b in enumerate(self.buckets):
            for p in b:
                self.bucket_of[p] = bi
        self.flat = [None] * len(self.buckets)   # persistent flat storage, allocated on first use
        self.handles = []
        self.enabled = True   # False: gradients stay local (the other network's phase; single-rank diagnostics)
        self.avg_op = None
        self.start()
        if self.world > 1:
            backend = dist.get_backend(group)
            self.avg_op = dist.ReduceOp.AVG if backend == "nccl" else None
            for p in params:
                self.handles.append(p.register_post_accumulate_grad_hook(self._hook))

    def start(self):
        """Forget the bookkeeping of any earlier backward pass (call before the backward of an update)."""
        self.pending = [len(b) for b in self.buckets]
        self.launched = [False] * len(self.buckets)
        self.inflight = []

    def _hook(self, p):
        if not self.enabled:
            return
        bi = self.bucket_of[p]
        self.pending[bi] -= 1
        if self.pending[bi] == 0 and not self.launched[bi]:
            self._launch(bi)

    def _launch(self, bi):
        ps = [p for p in self.buckets[bi] if p.grad is not None]
        self.launched[bi] = True
        if not ps:
            return
        n = sum(p.numel() for p in ps)
        flat = self.flat[bi]
        if flat is None or flat.numel() < n or flat.device != ps[0].grad.device:
            flat = self.flat[bi] = torch.empty(sum(p.numel() for p in self.buckets[bi]), dtype=torch.float32,
                                               device=ps[0].grad.device)
        views, off = [], 0
        for p in ps:
            views.append(flat[off:off + p.numel()].view_as(p.grad))
            off += p.numel()
        torch._foreach_copy_(views, [p.grad for p in ps])
        buf = flat[:n]
        if self.avg_op is not None:
            work = dist.all_reduce(buf, op=self.avg_op, group=self.group, async_op=True)
        else:
            buf.mul_(1.0 / self.world)
            work = dist.all_reduce(buf, group=self.group, async_op=True)
        self.inflight.append((ps, views, work))

    def finish(self):
        """Wait for all buckets (launching the ones backward did not complete) and write the
        averaged gradients back.  Call between backward() and optimizer.step()."""
        if self.world == 1 or not self.enabled:
            return
        for bi in range(len(self.buckets)):
            if not self.launched[bi] and self.pending[bi] < len(self.buckets[bi]):
                self._launch(bi)
        for ps, views, work in self.inflight:
            work.wait()
            torch._foreach_copy_([p.grad for p in ps], views)
        self.start()


class comm(object):
    def __init__(self, mode="env", backend=None):
        """mode 'env': RANK/WORLD_SIZE/LOCAL_RANK/MASTER_* from the environment (torchrun);
        mode 'dummy': single process, no process group (comm/distributed.py:58-66)."""
        self._local_rank = int(os.environ.get("LOCAL_RANK", 0))
        if mode != "dummy" and int(os.environ.get("WORLD_SIZE", "1")) > 1 and not dist.is_initialized():
            if backend is None:
                backend = "nccl" if torch.cuda.is_available() else "gloo"
            if backend == "nccl":
                torch.cuda.set_device(self._local_rank)
            dist.init_process_group(backend=backend)

    def size(self):
        return dist.get_world_size() if dist.is_initialized() else 1

    def rank(self):
        return dist.get_rank() if dist.is_initialized() else 0

    def local_rank(self):
        return self._local_rank

    def metric_average(self, val, name=None, op_name=None, device=None):
        """clone -> all_reduce(SUM) -> .item(); `op_name='average'` divides by size, the default
        (None) is a plain sum exactly like comm/distributed.py:12-34."""
        if not torch.is_tensor(val):
            val = torch.tensor(val, device=device)
        t = val.detach().clone().float()
        if dist.is_initialized():
            dist.all_reduce(t)
            if op_name == "average":
                t /= self.size()
        return t.item()

    def broadcast(self, tensor, root_rank=0, name=None):
        if dist.is_initialized():
            dist.broadcast(tensor, root_rank)
        return tensor

    def printr(self, msg, rank=0):
        if self.rank() == rank:
            print(msg, flush=True)

    def DistributedModel(self, model, bucket_mb=32):
        """Broadcast parameters/buffers from rank 0 and attach the bucketed gradient all-reduce as
        `model.grad_sync` (call `model.grad_sync.finish()` before optimizer.step())."""
        if dist.is_initialized():
            with torch.no_grad():
                for t in list(model.parameters()) + list(model.buffers()):
                    dist.broadcast(t, 0)
        model.grad_sync = BucketedAllReduce(model, bucket_mb)
        return model

    def DistributedOptimizer(self, optimizer, named_parameters=None, compression_name=None, op_name=None):
        return optimizer

    # ---- checkpoint interchange (comm/distributed.py:108-157; writer: train_gan.py:401-431) -------
    @staticmethod
    def _strip(sd):
        """The reference's key hack: apex/torch DDP prefix every key with `module.`."""
        return {k.replace("module.", ""): v for k, v in sd.items()}

    def init_training_state(self, model, optimizer, checkpoint_name, device_id):
        if (checkpoint_name is not None) and os.path.isfile(checkpoint_name):
            checkpoint = torch.load(checkpoint_name, map_location=device_id)
            optimizer.load_state_dict(checkpoint["optimizer"])
            model.load_state_dict(self._strip(checkpoint["model"]))
            return checkpoint["step"], checkpoint["epoch"]
        return 0, 0

    def init_gan_training_state(self, gmodel, dmodel, gopt, dopt, checkpoint_name, device_id):
        """Restart from a `.cpt` written by the reference trainer (or by save_gan_checkpoint): the
        parameter names are the compatibility contract, so the files interchange both ways."""
        if (checkpoint_name is not None) and os.path.isfile(checkpoint_name):
            checkpoint = torch.load(checkpoint_name, map_location=device_id)
            gopt.load_state_dict(checkpoint["g_opt"])
            dopt.load_state_dict(checkpoint["d_opt"])
            gmodel.load_state_dict(self._strip(checkpoint["generator"]))
            dmodel.load_state_dict(self._strip(checkpoint["discriminator"]))
            return checkpoint["step"], checkpoint["epoch"]
        return 0, 0

    def save_gan_checkpoint(self, path, step, epoch, generator, discriminator, g_opt, d_opt):
        """The dictionary train_gan.py:402-411 saves (`amp` is empty: opt level O0 has no loss-scaler state)."""
        if self.rank() == 0:
            torch.save({"step": step, "epoch": epoch, "generator": generator.state_dict(),
                        "discriminator": discriminator.state_dict(), "g_opt": g_opt.state_dict(),
                        "d_opt": d_opt.state_dict(), "amp": {}}, path)

