"""One IQ training step on the device: forward + losses + hand-written backward (+ gradient all-reduce) + clip + Adam.

Replaces what Lightning does around `TrainIQ.training_step` (train_iq.py:105-132): `loss.backward()`,
`gradient_clip_val=5` (train_iq.py:372), `Adam.step()` (train_iq.py:259-261) with the Noam learning rate of
`custom_optimizer` (train_iq.py:252-257) and the pre-training -> latent switch (train_iq.py:108-111).
The loss terms stay on the device (one 8-float vector); nothing synchronises with the host inside a step.
"""
import math
import os

import torch

from . import _lib, ops

HIGH_PRIORITY_STEP = os.environ.get("B200IQ_HP_STREAM", "1") == "1"
# B200IQ_PDL=1: programmatic dependent launch for the whole step (the next kernel's prologue runs under the previous
# kernel's tail).  Off by default: with the side streams filling the tails it is a net loss (see iq.py PDL_FORWARD).
PDL = os.environ.get("B200IQ_PDL", "3") == "1"


def noam_lr(step, hidden_dim, warmup=4000):
    """train_iq.py:252-257 (`lr(0) = 0`, peak at step == warmup)"""
    return math.sqrt(1.0 / hidden_dim) * min(math.sqrt(1.0 / (step + 1)), step * warmup ** -1.5)


class TrainStep:
    """owns the Adam moments and the per-step schedule for one `IQ` module on one GPU (one DP rank)"""

    def __init__(self, model, max_norm=5.0, betas=(0.9, 0.999), eps=1e-8, lr_fn=None, dp=None):
        self.model = model
        self.max_norm, self.betas, self.eps = max_norm, betas, eps
        flat = model.flat.data
        self.exp_avg = torch.zeros_like(flat)
        self.exp_avg_sq = torch.zeros_like(flat)
        self.hyper = torch.zeros(4, device=flat.device)
        self.normsq = torch.zeros(1, device=flat.device)
        self._hp = None
        self.iter = 0      # TrainIQ.iter (train_iq.py:43)
        self.kliter = 0    # TrainIQ.kliter (:44), counts steps since the latent switch
        self.t = 0         # Adam step count (never reset: train_iq.py:111 is a no-op)
        self.lr_fn = lr_fn or (lambda s: noam_lr(s, model.H))
        self.dp = dp       # parallel.GradReducer or None

    def set_hyper(self):
        lr = self.lr_fn(self.iter)
        t = self.t + 1
        # grad_scale: the data-parallel mean and the loss scale of the pass that just filled the gradient arena (fp16-plane path,
        # IQ.scale()) leave the gradients here
        scale = getattr(self.model, "_grad_scale", 1.0)
        ops.set_hyper(self.hyper, lr, 1.0 - self.betas[0] ** t, 1.0 - self.betas[1] ** t,
                      1.0 / ((self.dp.world_size if self.dp is not None else 1) * scale))

    def step(self, images, answers, response, target, eps=None):
        """returns the device tensor of loss terms (rec, img, kld, aux, count, gnorm_sq, ...) of this step"""
        if not HIGH_PRIORITY_STEP:
            with _lib.pinned_stream():
                return self._step(images, answers, response, target, eps)
        # the step's critical path runs on a high-priority stream, so that when one of its GEMMs and an off-path kernel on a
        # (default-priority) side stream are both pending, the block scheduler hands the SMs to the critical path first
        # and the side kernel fills the last wave.  Ordered against the caller's stream on both sides.
        cur = torch.cuda.current_stream()
        if self._hp is None:
            self._hp = torch.cuda.Stream(priority=-1)
        self._hp.wait_stream(cur)
        with torch.cuda.stream(self._hp), _lib.pinned_stream():
            terms = self._step(images, answers, response, target, eps)
        cur.wait_stream(self._hp)
        return terms

    def _step(self, images, answers, response, target, eps):
        prev = ops.set_pdl(PDL)
        try:
            return self._step_body(images, answers, response, target, eps)
        finally:
            ops.set_pdl(prev)

    def _step_body(self, images, answers, response, target, eps):
        m = self.model
        a = m.args
        if 0 <= getattr(a, "num_pretraining_steps", -1) <= self.iter and not m.latent_transformer:
            # train_iq.py:108-110 switches when iter == num_pretraining_steps; ">=" also covers a run resumed from a checkpoint
            # taken after the switch (the reference's checkpoint carries no mode flag)
            m.switch_GVT_train_mode(True)
        if self.dp is not None:
            self.dp.begin_step(m)
        terms = m.forward_losses(images, answers, response, target, kliter=self.kliter, eps=eps, backward=True)
        self.set_hyper()   # after the pass: its loss scale is part of grad_scale; only the Adam kernel below reads `hyper`
        g = m.grad_arena
        if self.dp is not None:
            self.dp.finish_step(m)
        ops.grad_sumsq(g, self.normsq)
        ops.adam_step(m.flat.data, g, self.exp_avg, self.exp_avg_sq, self.hyper, self.normsq, self.betas[0], self.betas[1],
                      self.eps, self.max_norm, zero_grad=True)
        self.iter += 1
        self.t += 1
        if m.latent_transformer:
            self.kliter += 1
        return terms
