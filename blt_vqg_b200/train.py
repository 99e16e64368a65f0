"""One IQ training step on the device: forward + losses + hand-written backward (+ gradient all-reduce) + clip + Adam.

Replaces what Lightning does around `TrainIQ.training_step` (train_iq.py:105-132): `loss.backward()`,
`gradient_clip_val=5` (train_iq.py:372), `Adam.step()` (train_iq.py:259-261) with the Noam learning rate of
`custom_optimizer` (train_iq.py:252-257) and the pre-training -> latent switch (train_iq.py:108-111).
The loss terms stay on the device (one 8-float vector); nothing synchronises with the host inside a step.

CUDA graph: a step is ~600 launches over up to five streams and costs ~12 ms of Python + ctypes to issue — as long as the
device needs; so for the transformer IQ the whole step (every stream fork / join, the NCCL buckets included) is captured once
per input shape and replayed: one launch per step.  What changes from step to step is read from device memory, because a
captured launch keeps its arguments: the per-step part of the Philox keys (`IQ.seed_dev`, bumped inside the graph), the annealed
KL coefficient (`IQ.kld_coef_dev`) and Adam's {lr, bias corrections, grad scale} (`hyper`) — the last two are written by two
tiny eager launches in front of each replay.  `B200IQ_GRAPH=0` (or `graph=False`) keeps the eager schedule.
"""
import math
import os

import torch

from . import _lib, ops

HIGH_PRIORITY_STEP = os.environ.get("B200IQ_HP_STREAM", "1") == "1"
# B200IQ_PDL=1: programmatic dependent launch for the whole step (the next kernel's prologue runs under the previous
# kernel's tail).  Off by default: with the side streams filling the tails it is a net loss (see iq.py PDL_FORWARD).
PDL = os.environ.get("B200IQ_PDL", "3") == "1"
GRAPH = os.environ.get("B200IQ_GRAPH", "1") == "1"
# B200IQ_GRAPH_PDL=1: programmatic dependent launch for every kernel of the captured step (programmatic edges in the graph: a kernel's
# prologue — barrier init, TMEM allocation, tensor-map fetch — runs under the tail of its predecessor); =2: only where the eager
# schedule uses it (the decoder's forward)
GRAPH_PDL = int(os.environ.get("B200IQ_GRAPH_PDL", "0"))
GRAPH_WARMUP = 2   # eager steps per input shape before the capture (workspaces, kernel attributes, NCCL channels are set up by then)


def noam_lr(step, hidden_dim, warmup=4000):
    """train_iq.py:252-257 (`lr(0) = 0`, peak at step == warmup)"""
    return math.sqrt(1.0 / hidden_dim) * min(math.sqrt(1.0 / (step + 1)), step * warmup ** -1.5)


class _Captured:
    """one captured step: static input buffers, the graph, its output"""

    def __init__(self):
        self.inputs, self.graph, self.terms, self.warm, self.launches, self.collectives = None, None, None, 0, 0, 0


class TrainStep:
    """owns the Adam moments and the per-step schedule for one `IQ` module on one GPU (one DP rank)"""

    def __init__(self, model, max_norm=5.0, betas=(0.9, 0.999), eps=1e-8, lr_fn=None, dp=None, graph=None):
        self.model = model
        self.max_norm, self.betas, self.eps = max_norm, betas, eps
        flat = model.flat.data
        self.exp_avg = torch.zeros_like(flat)
        self.exp_avg_sq = torch.zeros_like(flat)
        self.hyper = torch.zeros(4, device=flat.device)
        self.coefs = torch.zeros(4, device=flat.device)   # [0] = KL coefficient * loss scale of the step (graph mode)
        self.normsq = torch.zeros(1, device=flat.device)
        self._hp = None
        self.iter = 0      # TrainIQ.iter (train_iq.py:43)
        self.kliter = 0    # TrainIQ.kliter (:44), counts steps since the latent switch
        self.t = 0         # Adam step count (never reset: train_iq.py:111 is a no-op)
        self.lr_fn = lr_fn or (lambda s: noam_lr(s, model.H))
        self.dp = dp       # parallel.GradReducer or None
        # graphs need the models' device-resident per-step scalars (seed_dev / kld_coef_dev)
        self.graph = (GRAPH if graph is None else graph) and hasattr(model, "seed_dev") and flat.is_cuda
        self._captured = {}
        self._seed_ctr = None
        self.graph_replays = 0

    def set_hyper(self):
        lr = self.lr_fn(self.iter)
        t = self.t + 1
        # grad_scale: the data-parallel mean and the loss scale of the pass that fills the gradient arena (fp16-plane path,
        # IQ.scale()) leave the gradients here
        scale = getattr(self.model, "_grad_scale", 1.0)
        ops.set_hyper(self.hyper, lr, 1.0 - self.betas[0] ** t, 1.0 - self.betas[1] ** t,
                      1.0 / ((self.dp.world_size if self.dp is not None else 1) * scale))

    def _hp_stream(self):
        if self._hp is None:
            self._hp = torch.cuda.Stream(priority=-1) if HIGH_PRIORITY_STEP else torch.cuda.Stream()
        return self._hp

    def step(self, images, answers, response, target, eps=None):
        """returns the device tensor of loss terms (rec, img, kld, aux, count, gnorm_sq, ...) of this step"""
        m = self.model
        a = m.args
        if 0 <= getattr(a, "num_pretraining_steps", -1) <= self.iter and not m.latent_transformer:
            # train_iq.py:108-110 switches when iter == num_pretraining_steps; ">=" also covers a run resumed from a checkpoint
            # taken after the switch (the reference's checkpoint carries no mode flag)
            m.switch_GVT_train_mode(True)
        if self.graph and ops.GEMM_TRACE is None and not ops.exact_mode():
            terms = self._graph_step(images, answers, response, target, eps)
        elif not HIGH_PRIORITY_STEP:
            with _lib.pinned_stream():
                terms = self._step(images, answers, response, target, eps)
        else:
            # the step's critical path runs on a high-priority stream, so that when one of its GEMMs and an off-path kernel on a
            # (default-priority) side stream are both pending, the block scheduler hands the SMs to the critical path first
            # and the side kernel fills the last wave.  Ordered against the caller's stream on both sides.
            cur = torch.cuda.current_stream()
            hp = self._hp_stream()
            hp.wait_stream(cur)
            with torch.cuda.stream(hp), _lib.pinned_stream():
                terms = self._step(images, answers, response, target, eps)
            cur.wait_stream(hp)
        self.iter += 1
        self.t += 1
        if m.latent_transformer:
            self.kliter += 1
        return terms

    def _step(self, images, answers, response, target, eps):
        prev = ops.set_pdl(PDL)
        try:
            return self._step_body(images, answers, response, target, eps)
        finally:
            ops.set_pdl(prev)

    def _step_body(self, images, answers, response, target, eps, hyper_set=False):
        """the device work of one step (launches only; the counters are advanced by step())"""
        m = self.model
        if self.dp is not None:
            self.dp.begin_step(m)
        terms = m.forward_losses(images, answers, response, target, kliter=self.kliter, eps=eps, backward=True)
        g = m.grad_arena
        if self.dp is not None:
            self.dp.finish_step(m)
        if not hyper_set:
            self.set_hyper()   # after the pass: its loss scale is part of grad_scale; only the Adam kernel below reads `hyper`
        ops.grad_sumsq(g, self.normsq)
        ops.adam_step(m.flat.data, g, self.exp_avg, self.exp_avg_sq, self.hyper, self.normsq, self.betas[0], self.betas[1],
                      self.eps, self.max_norm, zero_grad=True)
        return terms

    # ------------------------------------------------------------------ CUDA-graph path
    def _static_inputs(self, xs):
        dev = self.model.flat.device
        out = []
        for x in xs:
            if x is None:
                out.append(None)
            else:
                dt = torch.float32 if x.is_floating_point() else torch.int32
                out.append(torch.empty(tuple(x.shape), device=dev, dtype=dt))
        return out

    def _graph_step(self, images, answers, response, target, eps):
        m = self.model
        xs = (images, answers, response if m.latent_transformer else None, target, eps)
        key = tuple((tuple(x.shape), x.is_floating_point()) if x is not None else None for x in xs) + (
            bool(m.latent_transformer), bool(m.training), id(getattr(m, "feature_cache", None)), getattr(m, "attn_dropout", 0.0),
            getattr(m, "relu_dropout", 0.0), getattr(m, "input_dropout_p", 0.0))
        st = self._captured.get(key)
        if st is None:
            st = self._captured[key] = _Captured()
            st.inputs = self._static_inputs(xs)
        if self._seed_ctr is None:
            # from here on the device word counts the stochastic passes (it starts where the Python counter stands)
            self._seed_ctr = torch.full((1,), int(m.step_idx), device=m.flat.device, dtype=torch.int64)
        m.seed_dev, m.kld_coef_dev = self._seed_ctr, self.coefs
        cur = torch.cuda.current_stream()
        hp = self._hp_stream()
        hp.wait_stream(cur)
        try:
            with torch.cuda.stream(hp), _lib.pinned_stream():
                for s, x in zip(st.inputs, xs):
                    if x is not None and s.data_ptr() != x.data_ptr():
                        s.copy_(x, non_blocking=True)
                # the two per-step scalars that come from the host: Adam's hyper-parameters and the annealed KL coefficient
                c_kld = m.loss_coefs(self.kliter)[2] if hasattr(m, "loss_coefs") else 0.0
                if st.warm > 0:
                    self.set_hyper()
                    ops.set_hyper(self.coefs, c_kld * m._grad_scale, 0.0, 0.0, 0.0)
                if st.graph is None and st.warm < GRAPH_WARMUP:
                    # eager warm-up on the static buffers.  The very first pass also finds the loss scale (known from the shapes)
                    if st.warm == 0:
                        ops.set_hyper(self.coefs, c_kld * m.scale(target.shape[0] * target.shape[1]), 0.0, 0.0, 0.0)
                    terms = self._device_step(st, hyper_set=st.warm > 0)
                    st.warm += 1
                elif st.graph is None:
                    terms = self._capture(st, hp)
                else:
                    st.graph.replay()
                    self.graph_replays += 1
                    ops.LAUNCHES += st.launches   # what the replay launched (bench.py reports gpu_launches)
                    if self.dp is not None:
                        self.dp.collectives += st.collectives
                    terms = st.terms
        finally:
            m.seed_dev, m.kld_coef_dev = None, None
        m.step_idx += 1   # mirrors the device counter (checkpoint.py reads it)
        cur.wait_stream(hp)
        return terms

    def _device_step(self, st, hyper_set):
        allowed, ops.PDL_ALLOWED = ops.PDL_ALLOWED, GRAPH_PDL > 0   # default: no programmatic edges inside the captured step
        prev = ops.set_pdl(GRAPH_PDL == 1)
        try:
            terms = self._step_body(st.inputs[0], st.inputs[1], st.inputs[2], st.inputs[3], st.inputs[4], hyper_set=hyper_set)
            self._seed_ctr.add_(1)
            return terms
        finally:
            ops.PDL_ALLOWED = allowed
            ops.set_pdl(prev)

    def _capture(self, st, hp):
        """capture one step on the (already current) high-priority stream; the first replay IS this step"""
        l0 = ops.LAUNCHES
        c0 = self.dp.collectives if self.dp is not None else 0
        g = torch.cuda.CUDAGraph()
        hp.synchronize()
        with torch.cuda.graph(g, stream=hp, capture_error_mode="thread_local"), _lib.pinned_stream():
            st.terms = self._device_step(st, hyper_set=True)
        st.launches = ops.LAUNCHES - l0
        ops.LAUNCHES = l0
        if self.dp is not None:
            st.collectives = self.dp.collectives - c0
        st.graph = g
        g.replay()
        self.graph_replays += 1
        ops.LAUNCHES += st.launches
        return st.terms
