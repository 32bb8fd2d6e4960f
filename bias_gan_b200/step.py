"""One G+D training step — the compute hot loop of src/deepCam/gpsro_train/train_gan.py:244-301,
on staged bf16 fields.

`lean=True` (default) drops work whose results the reference discards: the D phase runs G without
building a graph (the reference back-propagates d_loss into G and then zeroes those gradients,
SURVEY F8) and the G phase does not compute D's weight gradients (zeroed by the next
d_opt.zero_grad()).  Parameters, BN running statistics, losses and RNG draw order after the step
are identical to the reference-faithful schedule (`lean=False`).
"""
import torch


def gan_step(G, D, g_opt, d_opt, crit_gan, crit_reg, inputs, real, w_gan=1.0, w_reg=1.0, lean=True,
             train_discriminator=True, train_generator=True, warmup=False):
    """inputs / real: bf16 field tensors [N, H/2, W/2, 4C].  Returns (d_loss, g_loss) as 0-d device
    tensors (no host sync; comm.metric_average decides when to read them)."""
    d_loss = g_loss = None
    if train_discriminator:
        if lean:
            with torch.no_grad():
                fake = G.forward_field(inputs)
        else:
            fake = G.forward_field(inputs)
        logits_real, _ = D.forward_field(real)
        logits_fake, _ = D.forward_field(fake)
        d_loss = crit_gan.d_loss(logits_real, logits_fake)
        d_opt.zero_grad(set_to_none=True)
        d_loss.backward()
        if hasattr(D, "grad_sync"):
            D.grad_sync.finish()
        d_opt.step()
        del fake, logits_real, logits_fake
    if train_generator:
        fake = G.forward_field(inputs)
        if lean:
            for p in D.parameters():
                p.requires_grad_(False)
        try:
            logits_fake, _ = D.forward_field(fake)
            gan_loss = crit_gan.g_loss(logits_fake)
            reg_loss = crit_reg(fake, real)
            g_loss = reg_loss if warmup else w_gan * gan_loss + w_reg * reg_loss
            g_opt.zero_grad(set_to_none=True)
            g_loss.backward()
        finally:
            if lean:
                for p in D.parameters():
                    p.requires_grad_(True)
        if hasattr(G, "grad_sync"):
            G.grad_sync.finish()
        g_opt.step()
    return (d_loss.detach() if d_loss is not None else None,
            g_loss.detach() if g_loss is not None else None)
