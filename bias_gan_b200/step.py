"""One G+D training step — the compute hot loop of src/deepCam/gpsro_train/train_gan.py:244-301,
on staged bf16 fields.

`lean=True` (default) drops work whose results the reference discards: the D phase runs G without
building a graph (the reference back-propagates d_loss into G and then zeroes those gradients,
SURVEY F8) and the G phase does not compute D's weight gradients (zeroed by the next
d_opt.zero_grad()).  Parameters, BN running statistics, losses and RNG draw order after the step
are identical to the reference-faithful schedule (`lean=False`).
"""
import contextlib
import os

import torch

from . import layouts as L

# one bgr_weights_prepare launch per network and phase instead of per-layer weight copies (A/B switch)
PACK_WEIGHTS = os.environ.get("BGR_PACK_WEIGHTS", "1") != "0"


@contextlib.contextmanager
def _sync_only(stepped, other):
    """The backward pass of `stepped`'s update: its gradient buckets start clean, and `other`'s hooks stay
    silent (the reference schedule back-propagates d_loss into G and g_loss into D; those gradients are
    discarded, so they must neither be all-reduced nor leave bucket state behind)."""
    so, oo = getattr(stepped, "grad_sync", None), getattr(other, "grad_sync", None)
    if so is not None:
        so.start()
    was = oo.enabled if oo is not None else None
    if oo is not None:
        oo.enabled = False
    try:
        yield
    finally:
        if oo is not None:
            oo.enabled = was


def gan_step(G, D, g_opt, d_opt, crit_gan, crit_reg, inputs, real, w_gan=1.0, w_reg=1.0, w_gp=10.0, lean=True,
             train_discriminator=True, train_generator=True, warmup=False, masks=None):
    """inputs / real: bf16 field tensors [N, H/2, W/2, 4C]; masks (optional, --enable_masks): per-element
    weights in the same layout, handed to the regression criterion as train_gan.py:278-279 does.  Returns
    (d_loss, g_loss) as 0-d device tensors (no host sync; comm.metric_average decides when to read them)."""
    d_loss = g_loss = None
    prep = L.prepare_weights if (PACK_WEIGHTS and inputs.is_cuda) else (lambda m: None)
    if train_discriminator:
        prep(G)
        prep(D)
        if lean:
            with torch.no_grad():
                fake = G.forward_field(inputs)
        else:
            fake = G.forward_field(inputs)
        logits_real, _ = D.forward_field(real)
        logits_fake, _ = D.forward_field(fake)
        d_loss = crit_gan.d_loss(logits_real, logits_fake)
        if crit_gan.mode == "Wasserstein":   # train_gan.py:257-258
            from .deeplab_gan import gradient_penalty
            d_loss = d_loss + w_gp * gradient_penalty(D, fake, real, eta=crit_gan.static_eta())
        d_opt.zero_grad(set_to_none=True)
        with _sync_only(D, G):
            d_loss.backward()
        if hasattr(D, "grad_sync"):
            D.grad_sync.finish()
        d_opt.step()
        L.invalidate_cache(D)
        del fake, logits_real, logits_fake
    if train_generator:
        prep(G)
        prep(D)      # D changed in the critic phase: rebuilt once here instead of layer by layer
        fake = G.forward_field(inputs)
        if lean:
            for p in D.parameters():
                p.requires_grad_(False)
        try:
            logits_fake, _ = D.forward_field(fake)
            gan_loss = crit_gan.g_loss(logits_fake)
            reg_loss = crit_reg(fake, real, masks) if masks is not None else crit_reg(fake, real)
            g_loss = reg_loss if warmup else w_gan * gan_loss + w_reg * reg_loss
            g_opt.zero_grad(set_to_none=True)
            with _sync_only(G, D):
                g_loss.backward()
        finally:
            if lean:
                for p in D.parameters():
                    p.requires_grad_(True)
        if hasattr(G, "grad_sync"):
            G.grad_sync.finish()
        g_opt.step()
        L.invalidate_cache(G)
    return (d_loss.detach() if d_loss is not None else None,
            g_loss.detach() if g_loss is not None else None)


class GraphedGanStep:
    """gan_step captured once into a CUDA graph and replayed: the ~3400 kernel launches of a step
    (and PyTorch's autograd bookkeeping between them) collapse into one graph launch, so the GPU is
    never waiting for the host.  Inputs are copied (or staged directly, see `inputs`/`real`) into
    static field buffers; the adversarial labels are drawn on the host before each replay
    (GANLoss.draw_labels) so the captured region has no host RNG and no host->device copy.

    Optimizers must be capturable (torch.optim.AdamW(..., capturable=True)).  Model, BatchNorm and
    optimizer state are snapshotted before the warm-up iterations that capture needs and restored
    in place afterwards, so constructing this object does not advance training.
    """

    def __init__(self, G, D, g_opt, d_opt, crit_gan, crit_reg, field_shape, device, warmup_iters=2, **step_kw):
        self.G, self.D, self.g_opt, self.d_opt = G, D, g_opt, d_opt
        self.crit_gan, self.crit_reg, self.step_kw = crit_gan, crit_reg, step_kw
        self.inputs = torch.zeros(field_shape, dtype=torch.bfloat16, device=device)
        self.real = torch.zeros(field_shape, dtype=torch.bfloat16, device=device)
        self.masks = torch.ones(field_shape, dtype=torch.bfloat16, device=device) if step_kw.pop("use_masks", False) else None
        self.noisy = getattr(G, "noise_dimensions", 0) > 0
        cpu_rng = torch.get_rng_state()
        crit_gan.draw_labels()
        if self.noisy:
            G.static_noise(True)          # slots are created (and filled once) by the warm-up forwards
        # state that the warm-up steps touch: parameters, buffers, optimizer moments
        tensors = [t for m in (G, D) for t in list(m.parameters()) + list(m.buffers())]
        side = torch.cuda.Stream(device=device)
        side.wait_stream(torch.cuda.current_stream(device))
        def opt_tensors():
            return [t for o in (g_opt, d_opt) for st in o.state.values() for t in st.values() if torch.is_tensor(t)]

        with torch.cuda.stream(side):
            snap = [t.detach().clone() for t in tensors]
            opt_before = {id(t): t.detach().clone() for t in opt_tensors()}
            opt_snap = None
            for i in range(max(1, warmup_iters)):
                self._run()
                if i == 0:  # lazily created optimizer state: its value before the first step is zero
                    opt_snap = [(t, opt_before.get(id(t), None)) for t in opt_tensors()]
                    opt_snap = [(t, torch.zeros_like(t) if s0 is None else s0) for t, s0 in opt_snap]
            L.invalidate_cache()
        torch.set_rng_state(cpu_rng)      # the capture-time draws must not advance the label / noise stream
        torch.cuda.current_stream(device).wait_stream(side)
        torch.cuda.synchronize(device)
        from . import _lib
        lib = _lib.lib()
        lib.bgr_launch_count.restype = __import__("ctypes").c_ulonglong
        c0 = lib.bgr_launch_count()
        self.graph = torch.cuda.CUDAGraph()
        # BGR_GRAPH_PRIORITY=1 captures the critical chain on a HIGH-priority stream while the weight-gradient side
        # stream (functional._side_stream) keeps the default priority, so that at replay the block scheduler serves
        # the forward / data-gradient chain first.  Measured on B200 and NOT adopted (profiles/r02_sched_ab.txt):
        # 127.6 vs 126.3 ms/step -- the deferred weight gradients pile up in front of the optimizer step.
        cap = None
        if os.environ.get("BGR_GRAPH_PRIORITY", "0") == "1":
            cap = torch.cuda.Stream(device=device, priority=-1)
        with torch.cuda.graph(self.graph, stream=cap):
            self.d_loss, self.g_loss = self._run()
        self.launches_per_step = int(lib.bgr_launch_count() - c0)   # kernels of libbgr.so in the graph
        L.invalidate_cache()
        with torch.no_grad():
            for t, s0 in zip(tensors, snap):
                t.copy_(s0)
            for t, z in opt_snap:
                t.copy_(z)
        torch.cuda.synchronize(device)

    def _run(self):
        if self.noisy:
            self.G.rewind_noise()
        return gan_step(self.G, self.D, self.g_opt, self.d_opt, self.crit_gan, self.crit_reg, self.inputs, self.real,
                        masks=self.masks, **self.step_kw)

    def __call__(self, inputs=None, real=None, masks=None):
        """One training step.  inputs/real(/masks): bf16 fields to copy into the static buffers; pass None
        when the caller has staged straight into `self.inputs` / `self.real` (/ `self.masks`, which exists
        when the step was built with use_masks=True)."""
        if inputs is not None:
            self.inputs.copy_(inputs)
        if real is not None:
            self.real.copy_(real)
        if masks is not None:
            self.masks.copy_(masks)
        # host RNG in the reference's order: [G noise] -> label_fake, label_real, swap -> [G noise]
        if self.noisy:
            self.G.redraw_noise(0, 1)
        self.crit_gan.draw_labels()
        if self.noisy:
            self.G.redraw_noise(1, None)
        self.graph.replay()
        return self.d_loss, self.g_loss
