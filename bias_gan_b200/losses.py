"""GANLoss with the reference's interface (src/deepCam/utils/losses.py:129-172) and the fused
reconstruction loss.  The adversarial terms act on [B,1] logits (a handful of scalars, plain torch);
label noise and the 5 % swap are drawn on the CPU generator in the reference's order so seeded
runs reproduce the same labels.  The reconstruction loss over the full fields is the fused
loss-and-gradient kernel (functional.l1_loss).
"""
import torch
import torch.nn as nn

from . import functional as F_


class GANLoss:
    def __init__(self, mode, batch_size, device):
        self.mode = mode
        self.batch_size = batch_size
        self.device = device
        if mode == "ModifiedMinMax":
            self.criterion = nn.BCEWithLogitsLoss()
            self.label_real = torch.ones((batch_size, 1)).to(device)
            self.label_fake = torch.zeros((batch_size, 1)).to(device)
            self.dist_real = torch.distributions.uniform.Uniform(0.8, 1.0)
            self.dist_fake = torch.distributions.uniform.Uniform(0.0, 0.2)
            self.dist_swap = torch.distributions.uniform.Uniform(0.0, 1.0)
        elif mode == "Wasserstein":
            self.dist_eta = torch.distributions.uniform.Uniform(0.0, 1.0)   # gradient_penalty's draw
            self._eta = None
        else:
            raise NotImplementedError("Error, {} loss not implemented".format(mode))
        self._static = None   # (label_for_fake, label_for_real) device buffers once draw_labels() is used
        self._copied = None   # CUDA event: the last host->device copy out of the pinned staging buffers

    def _staging_free(self):
        """The pinned staging buffers are rewritten every step: wait until the previous step's asynchronous
        host->device copies have read them (a host running ahead of graph replays would otherwise hand
        step i the labels of step i+1)."""
        if self._copied is not None:
            self._copied.synchronize()

    def _staging_used(self):
        if torch.cuda.is_available() and torch.device(self.device).type == "cuda":
            self._copied = torch.cuda.Event()
            self._copied.record()

    def draw_labels(self):
        """Draw this step's noisy labels and the 5 % swap on the CPU generator — the same three
        draws in the same order as d_loss (losses.py:146-153) — into persistent device buffers.
        After the first call d_loss() reads those buffers instead of drawing, which keeps host RNG
        and host->device copies out of a captured CUDA graph (step.GraphedGanStep)."""
        if self.mode == "Wasserstein":   # the critic phase's only host draw: gradient_penalty's eta
            eta = self.dist_eta.rsample((self.batch_size, 1, 1, 1))
            if self._eta is None:
                pin = torch.cuda.is_available() and torch.device(self.device).type == "cuda"
                self._eta_host = torch.empty(eta.shape, pin_memory=pin)
                self._eta = torch.empty(eta.shape, device=self.device)
            self._staging_free()
            self._eta_host.copy_(eta)
            self._eta.copy_(self._eta_host, non_blocking=True)
            self._staging_used()
            return
        label_fake = self.dist_fake.rsample(self.label_fake.shape)
        label_real = self.dist_real.rsample(self.label_real.shape)
        if self.dist_swap.sample() < 0.05:
            label_fake, label_real = label_real, label_fake
        if self._static is None:
            pin = torch.cuda.is_available() and torch.device(self.device).type == "cuda"
            self._host = tuple(torch.empty(self.label_fake.shape, pin_memory=pin) for _ in range(2))
            self._static = tuple(torch.empty_like(self.label_fake) for _ in range(2))
        self._staging_free()
        for host, dev, val in zip(self._host, self._static, (label_fake, label_real)):
            host.copy_(val)
            dev.copy_(host, non_blocking=True)
        self._staging_used()

    def d_loss(self, logits_real, logits_fake):
        if self.mode == "Wasserstein":
            return torch.mean(logits_fake - logits_real)
        if self._static is not None:
            label_fake, label_real = self._static
            return 0.5 * (self.criterion(logits_fake, label_fake) + self.criterion(logits_real, label_real))
        label_fake = self.dist_fake.rsample(self.label_fake.shape).to(self.device, non_blocking=True)
        label_real = self.dist_real.rsample(self.label_real.shape).to(self.device, non_blocking=True)
        if self.dist_swap.sample() < 0.05:
            label_fake, label_real = label_real, label_fake
        return 0.5 * (self.criterion(logits_fake, label_fake) + self.criterion(logits_real, label_real))

    def static_eta(self):
        """The pre-drawn interpolation weights (None unless draw_labels() is in use)."""
        return getattr(self, "_eta", None)

    def g_loss(self, logits_fake):
        if self.mode == "Wasserstein":
            return -1.0 * torch.mean(logits_fake)
        return self.criterion(logits_fake, self.label_real)


class L1LossField(nn.Module):
    """nn.L1Loss (train_gan.py:146) on two fields in the same (space-to-depth bf16) layout.
    elements_per_sample = C*H*W of the real tensors when the fields carry zero padding (odd sizes, channel
    counts padded to 8): the padding contributes |0 - 0| to the sum and must not count in the mean."""

    def __init__(self, elements_per_sample=None):
        super().__init__()
        self.elements_per_sample = elements_per_sample

    def forward(self, fake_field, real_field):
        if self.elements_per_sample is None:
            return F_.l1_loss(fake_field, real_field)
        return F_.regression_loss(fake_field, real_field, None, 0, 1.0, self.elements_per_sample * fake_field.shape[0])


class RegressionLossField(nn.Module):
    """The reconstruction-loss choices of the training script (train_gan.py:141-152) on bf16 fields:
    kind 'l1' = nn.L1Loss, 'smooth_l1' = nn.SmoothL1Loss, 'l2' = nn.MSELoss; called with a third argument
    (the masks of --enable_masks, train_gan.py:278-279) it is utils/losses.py:101-112 L1LossWeighted
    (normalize=False: mean(l * w); normalize=True: sum(l * w) / (sum(w) + eps)) resp. L2LossWeighted.
    `count` = number of real elements when the fields carry zero padding (odd-sized inputs)."""
    MODES = {"l1": 0, "smooth_l1": 1, "l2": 2}

    def __init__(self, kind="l1", normalize=False, eps=1e-8, beta=1.0):
        super().__init__()
        if kind not in self.MODES:
            raise NotImplementedError("Error, loss {} not implemented.".format(kind))
        self.kind, self.normalize, self.eps, self.beta = kind, normalize, eps, beta

    def forward(self, fake_field, real_field, masks_field=None, count=None):
        return F_.regression_loss(fake_field, real_field, masks_field, self.MODES[self.kind], self.beta, count,
                                  self.normalize, self.eps)


def L1LossWeighted(normalize=False, eps=1e-8, smooth=False):
    """utils/losses.py:101-112 by name."""
    return RegressionLossField("smooth_l1" if smooth else "l1", normalize, eps)


def L2LossWeighted(normalize=False, eps=1e-8):
    return RegressionLossField("l2", normalize, eps)
