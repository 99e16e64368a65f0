#!/usr/bin/env python
"""bench.py — IQ train samples/sec (BASELINE.json metric) on N B200s of one node, plus the reference arm.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--workload transformer|lstm] [--impl ours|reference]
  torchrun --nproc-per-node N bench.py --gpus N ...      (one rank per GPU, NCCL)

A step = the whole hot path on one batch: image head -> encoders -> latent -> decoder -> fused vocab-projection+CE and
the other loss terms -> full hand-written backward -> (gradient all-reduce) -> global-norm clip + Adam.  The frozen CNN
trunk is excluded: inputs are pooled (B, FEAT) features.
Keys: `value` = device-timed throughput with inputs resident in HBM; `e2e` = the same through the public call with
pinned HOST buffers (H2D of the batch and D2H of the loss terms inside the timed region); `roofline` = the dominant
kernel (tcgen05 GEMM) timed with CUDA events on the launching stream; `cpu_baseline` = the oracle port of the reference
step on the host cores (bounded sample); `--impl reference` prints that CPU arm as its own line.
"""
import argparse
import json
import os
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

import torch  # noqa: E402

DTYPES = {"transformer": "f16 (fp16 operand planes: 3-term hi/lo forward products, single-pass backward under a loss scale; fp32 accumulate, fp32 master weights)",
          "lstm": "f16 (fp16 operand planes like the transformer; the persistent recurrence is a 3-term fp16 mma.sync product over hi / lo halves; fp32 accumulate, fp32 master weights)"}
METRIC = "iq_train_samples_per_sec"
UNIT = "samples/s"


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default=os.environ.get("B200IQ_WORKLOAD", "transformer"), choices=["transformer", "lstm"])
    ap.add_argument("--batch", type=int, default=0, help="per-GPU batch (default: the workload's)")
    ap.add_argument("--config", default="", help="other BASELINE.json configs as parity / sizing runs: transformer cfg1|cfg3|cfg5, lstm cfg1|cfg2|cfg4")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-roofline", action="store_true")
    ap.add_argument("--no-eager-baseline", action="store_true", help="skip the stock-PyTorch-eager-on-B200 timing of the oracle port")
    ap.add_argument("--no-secondary", action="store_true", help="skip the LSTM (configs[1]) measurement beside the transformer headline")
    ap.add_argument("--cpu-batch", type=int, default=64, help="bounded CPU sample of the cpu_baseline leg: batch size of the reference step")
    ap.add_argument("--ref-batch", type=int, default=0, help="--impl reference: batch (default: the workload's own, cut only if the time budget demands)")
    return ap.parse_args()


def workload_config(args):
    from blt_vqg_b200 import synth
    if args.workload == "transformer":
        cfg = dict(synth.CONFIGS[args.config or "cfg3"])   # BASELINE.json configs[2]: 6 layers, d_model 512, 8 heads, T=20, batch 512
        name = "IQ transformer decoder (BLT-VQG, latent mode), %d layers, d_model %d, %d heads, E=%d, F=%d, V=%d, T=%d, S=%d, R=%d, FEAT=%d" % (
            cfg["NL"], cfg["H"], cfg["NH"], cfg["E"], cfg["F"], cfg["V"], cfg["T"], cfg["S"], cfg["T"] + 1, cfg["FEAT"])
    else:
        cfg = dict(synth.LSTM_CONFIGS[args.config or "cfg2"])  # BASELINE.json configs[1]: GloVe-300 LSTM, V=15000, batch 256
        name = "IQ LSTM decoder%s, H=%d, Z=%d, V=%d, T=%d, FEAT=%d" % (" full size (GloVe 300-d => H=300)" if cfg["H"] == 300 else "", cfg["H"], cfg["Z"], cfg["V"], cfg["T"], cfg["FEAT"])
    if args.batch:
        cfg["B"] = args.batch
    return cfg, name


# ------------------------------------------------------------------ clocks during the timed region
class ClockSampler(threading.Thread):
    def __init__(self, index=0, period=0.1):
        super().__init__(daemon=True)
        self.index, self.period = index, period
        self.samples, self.reasons, self.max_mhz = [], set(), None
        self._stop_evt = threading.Event()
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
        except Exception:
            self.nv = None

    def run(self):
        if self.nv is None:
            return
        nv = self.nv
        names = {
            getattr(nv, "nvmlClocksEventReasonHwSlowdown", 0x8): "hw_slowdown",
            getattr(nv, "nvmlClocksEventReasonHwThermalSlowdown", 0x40): "hw_thermal_slowdown",
            getattr(nv, "nvmlClocksEventReasonSwThermalSlowdown", 0x20): "sw_thermal_slowdown",
            getattr(nv, "nvmlClocksEventReasonSwPowerCap", 0x4): "sw_power_cap",
        }
        while not self._stop_evt.is_set():
            try:
                self.samples.append(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM))
                try:
                    r = nv.nvmlDeviceGetCurrentClocksEventReasons(self.h)
                except Exception:
                    r = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                for bit, nm in names.items():
                    if r & bit:
                        self.reasons.add(nm)
            except Exception:
                pass
            time.sleep(self.period)

    def stop(self):
        self._stop_evt.set()
        self.join(timeout=2)
        s = sorted(self.samples)
        return {"sm_mhz": s[len(s) // 2] if s else None, "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons),
                "samples": len(s)}


# ------------------------------------------------------------------ the reference arm (the reference's OWN modules)
def _median(xs):
    xs = sorted(xs)
    return xs[len(xs) // 2] if len(xs) % 2 else 0.5 * (xs[len(xs) // 2 - 1] + xs[len(xs) // 2])


def reference_available():
    from oracle import refload
    return refload.available()


def build_reference_step(workload, cfg, device):
    """(params, step_fn(batch dict, step index) -> loss) over the UNMODIFIED reference modules from baseline/_ref (or
    /root/reference): transformer = models.IQ (trunk cut at pooled features, its own initialisers, dropout ON, eps from the CPU
    generator: stock behaviour) + TrainIQ.calculate_losses verbatim (train_iq.py:81-103); LSTM = the upstream glue over the
    reference's DecoderRNN / EncoderRNN / MLP classes (oracle/gen_golden_lstm.py RefLSTMIQ; the glue itself is absent from the
    tree).  Nothing of blt_vqg_b200's model code is on this path (synth only makes the synthetic batch)."""
    import torch.nn as nn
    import torch.nn.functional as F
    from types import SimpleNamespace
    from blt_vqg_b200 import synth
    from oracle import refload
    if workload == "transformer":
        rargs = synth.args_for(cfg, device=device)
        m = refload.build_reference_iq(synth.SynthVocab(cfg["V"]), rargs, True, seed=0, dropout=True).to(device).train()
        calc = refload.import_calculate_losses()
        self_ = SimpleNamespace(criterion=nn.CrossEntropyLoss(ignore_index=synth.PAD), image_recon_criterion=nn.MSELoss(),
                                latent_transformer=True, kliter=7000, args=rargs)

        def step_fn(b, step):
            output, z_logit, kld, (img, recon) = m(b["feats"], b["answers"], b["posteriors"], b["questions"])
            return calc(self_, output, (img, recon), kld, z_logit, b["questions"])[0]
    else:
        from oracle import lstm_oracle
        from oracle.gen_golden_lstm import RefLSTMIQ
        torch.manual_seed(0)
        m = RefLSTMIQ(cfg, synth.SOQ, synth.EOS)
        m.decoder.input_dropout.p = 0.3   # old_train_iq.py:539-564 defaults (dropout 0.3 / 0.3)
        m.answer_encoder.input_dropout.p = 0.3
        m = m.to(device).train()
        lam = lstm_oracle.LAMBDAS

        def step_fn(b, step):
            order = torch.argsort(b["alengths"], descending=True, stable=True)   # pack_padded_sequence wants sorted lengths
            b = {k: v[order].contiguous() for k, v in b.items()}
            img = m.encoder_cnn(b["feats"])
            ans = m.encode_answers(b["answers"], b["alengths"].tolist())
            mus, logvars = m.encode_into_z(img, ans)
            zs = torch.randn(mus.shape).to(mus.device) * torch.exp(0.5 * logvars) + mus
            outputs, _, _ = m.decode_questions(img, zs, questions=b["questions"], teacher_forcing_ratio=1.0)
            logp = torch.cat([o.unsqueeze(1) for o in outputs], dim=1)
            gen = F.nll_loss(logp.reshape(-1, cfg["V"]), b["questions"][:, 1:].reshape(-1), ignore_index=synth.PAD)
            kl = lstm_oracle.gaussian_kl_loss(mus, logvars)
            mus2, logvars2 = m.encode_into_z(img.detach(), ans.detach())
            z2 = torch.randn(mus2.shape).to(mus2.device) * torch.exp(0.5 * logvars2) + mus2
            rec_i = F.mse_loss(m.image_reconstructor(z2), img.detach())
            rec_a = F.mse_loss(m.answer_reconstructor(z2), ans.detach())
            return lam["lambda_gen"] * gen + lam["lambda_z"] * kl + lam["lambda_i"] * rec_i + lam["lambda_a"] * rec_a
    return [p for p in m.parameters() if p.requires_grad], step_fn


def build_port_step(workload, cfg, device):
    """fallback when no reference tree is reachable: the oracle restatement (kind "port")"""
    from blt_vqg_b200 import synth
    if workload == "transformer":
        from oracle import iq_oracle
        from blt_vqg_b200.iq import IQ
        torch.manual_seed(0)
        m0 = IQ(True, synth.SynthVocab(cfg["V"]), synth.args_for(cfg))
        P = {k: v.clone().to(device) for k, v in m0.state_dict().items()}
        params = [P[n].requires_grad_(True) for n in m0.arena.names()]
        ocfg = dict(num_layers=cfg["NL"], num_heads=cfg["NH"], latent_dim=cfg["L"], image_recon_lambda=0.1, kl_ceiling=0.5,
                    aux_ceiling=1.0, full_kl_step=15000)
        bn = (P["encoder_cnn.bn.running_mean"], P["encoder_cnn.bn.running_var"])

        def step_fn(b, step):
            eps = torch.randn(b["feats"].shape[0], cfg["L"]).to(device)
            out, zl, kld, img, recon = iq_oracle.iq_forward(P, ocfg, b["feats"], b["answers"], b["posteriors"], b["questions"],
                                                            eps=eps, latent_mode=True, training=True, drop=(0.1, 0.1), bn_state=bn)
            return iq_oracle.iq_losses(ocfg, out, zl, kld, img, recon, b["questions"], step, True)[0]
        return params, step_fn
    from oracle import lstm_oracle
    return lstm_oracle.make_training_closure(cfg)


def reference_steps(workload, cfg, steps, warmup, batch, device="cpu", mode="fp32", budget_s=None):
    """the reference training step — forward + losses + backward + clip_grad_norm_(5) + Adam (train_iq.py:259-261,372), fp32,
    stock PyTorch — on `device`.  device="cpu": all host threads.  device="cuda": the same modules as stock PyTorch eager on the
    B200 (the reference ships no kernels of its own, eager PyTorch IS its Blackwell path), mode "fp32" (TF32
    off) | "tf32" | "bf16" (autocast).  budget_s: if the first step predicts that warmup + steps would not fit, the batch is cut
    (multiples of 32) — the batch actually run is returned.  Returns a dict (value samples/s from the median step, min, cores, ...)."""
    from blt_vqg_b200 import synth
    torch.set_num_threads(os.cpu_count())
    cores = torch.get_num_threads()
    on_gpu = device != "cpu"
    if on_gpu:
        torch.backends.cuda.matmul.allow_tf32 = mode != "fp32"
        torch.backends.cudnn.allow_tf32 = mode != "fp32"
    kind = "reference" if reference_available() else "port"
    params, step_fn = (build_reference_step if kind == "reference" else build_port_step)(workload, cfg, device)
    opt = torch.optim.Adam(params, lr=1e-4)

    def one(s, B):
        b = synth.make_batch(cfg, seed=1234 + s, B=B)
        if on_gpu:
            b = {k: v.to(device) for k, v in b.items()}
            torch.cuda.synchronize()
        t0 = time.perf_counter()
        if mode == "bf16":
            with torch.autocast(device_type="cuda", dtype=torch.bfloat16):
                loss = step_fn(b, s)
        else:
            loss = step_fn(b, s)
        opt.zero_grad()
        loss.backward()
        torch.nn.utils.clip_grad_norm_(params, 5.0)
        opt.step()
        if on_gpu:
            torch.cuda.synchronize()
        return time.perf_counter() - t0

    times, s = [], 0
    if warmup > 0:
        t_first = one(s, batch)
        s += 1
        if budget_s is not None and t_first * (warmup + steps) > budget_s and batch > 32:
            batch = max(32, int(batch * budget_s / (t_first * (warmup + steps))) // 32 * 32)
        for _ in range(warmup - 1):
            one(s, batch)
            s += 1
    for _ in range(steps):
        times.append(one(s, batch))
        s += 1
    med = _median(times)
    return {"value": batch / med, "min_step_value": batch / max(times), "max_step_value": batch / min(times), "ms_per_step": 1e3 * med,
            "ms_per_step_min": 1e3 * min(times), "cores": cores, "kind": kind, "batch": batch, "steps": len(times), "warmup": warmup}


# ------------------------------------------------------------------ our arm
def build_ours(args, cfg, dp):
    from blt_vqg_b200 import _lib, synth
    from blt_vqg_b200.train import TrainStep
    _lib.require_device()
    torch.manual_seed(0)
    if args.workload == "transformer":
        from blt_vqg_b200.iq import IQ
        model = IQ(True, synth.SynthVocab(cfg["V"]), synth.args_for(cfg)).cuda().train()
    else:
        from blt_vqg_b200.lstm_iq import LSTMIQ
        model = LSTMIQ.from_config(cfg).cuda().train()
    # transformer: Noam schedule (train_iq.py:252-257) at its peak; LSTM: constant Adam lr (old_train_iq.py:358-359)
    ts = TrainStep(model, dp=dp, lr_fn=None if args.workload == "transformer" else (lambda s: 1e-4))
    ts.iter = 4000
    return model, ts


def host_batches(cfg, n, rank):
    """n distinct synthetic batches in pinned host memory, already in the kernels' dtypes (fp32 features, int32 ids)"""
    from blt_vqg_b200 import synth
    out = []
    for i in range(n):
        b = synth.make_batch(cfg, seed=1234 + 17 * i, rank=rank)
        out.append({k: (v.float() if v.is_floating_point() else v.int()).contiguous().pin_memory() for k, v in b.items()})
    return out


def step_inputs(workload, b):
    if workload == "transformer":
        return b["feats"], b["answers"], b["posteriors"], b["questions"]
    return b["feats"], b["answers"], b["alengths"], b["questions"]


def timed_region(fn, steps, dist_on):
    import torch.distributed as dist
    if dist_on:
        dist.barrier()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(steps):
        fn(i)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1)
    if dist_on:
        t = torch.tensor([ms], device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t)
    return ms


def kernel_rooflines(ts, model, args, dev_batches, peaks):
    """CUDA events on the launching stream around every launch of two extra EAGER steps (after the timed region; the CUDA-graph
    replay and the side streams are off for them, so one kernel runs at a time): the GEMMs against the measured tensor peak, the
    streaming kernels against the measured HBM peak, each with the algorithmic flops / bytes of its arguments"""
    from blt_vqg_b200 import iq as iq_mod, ops, transformer as tf_mod
    saved = (tf_mod.CONCURRENT, tf_mod.CONCURRENT_DW, iq_mod.CONCURRENT_ENCODERS)
    tf_mod.CONCURRENT = tf_mod.CONCURRENT_DW = iq_mod.CONCURRENT_ENCODERS = False
    ops.GEMM_TRACE, ops.KERNEL_TRACE = [], []
    step_t0, step_t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    try:
        # the eager issue of an instrumented step (every launch through ctypes plus two event records) is slower than the device
        # runs it; with an empty queue an event pair would then also span the host's gap in front of its kernel.  ~25 ms of
        # spinning first lets the host run ahead, so the launches of both steps wait in the stream behind their start events
        torch.cuda._sleep(int(25e-3 * 1.9e9))
        step_t0.record()
        for i in range(2):
            ts.step(*step_inputs(args.workload, dev_batches[i % len(dev_batches)]))
        step_t1.record()
        torch.cuda.synchronize()
    finally:
        tf_mod.CONCURRENT, tf_mod.CONCURRENT_DW, iq_mod.CONCURRENT_ENCODERS = saved
    trace, ktrace = ops.GEMM_TRACE, ops.KERNEL_TRACE
    ops.GEMM_TRACE, ops.KERNEL_TRACE = None, None
    step_ms = step_t0.elapsed_time(step_t1) / 2
    tot_ms = sum(e0.elapsed_time(e1) for e0, e1, *_ in trace)
    tot_fl = sum(t[2] for t in trace)
    hw_fl = sum(t[2] * (t[4][5] if len(t[4]) > 5 else 1) for t in trace)
    by_shape = {}
    for e0, e1, fl, by, tag, *_ in trace:
        d = by_shape.setdefault(tag, [0, 0.0, 0.0, 0.0])
        d[0] += 1
        d[1] += e0.elapsed_time(e1)
        d[2] += fl
        d[3] += by
    top = sorted(by_shape.items(), key=lambda kv: -kv[1][1])[:int(os.environ.get("B200IQ_GEMM_ROWS", "12"))]
    detail = [{"M": k[0], "N": k[1], "K": k[2], "flags": k[3], "split_k": k[4], "terms": k[5] if len(k) > 5 else "tf32", "launches": v[0], "ms": round(v[1], 4),
               "tflops": round(v[2] / v[1] * 1e-9, 1) if v[1] > 0 else None,
               "operand_gbs": round(v[3] / v[1] * 1e-6, 1) if v[1] > 0 else None} for k, v in top]
    peak = peaks.get("bf16_tflops_sustained") or 1400.0
    hbm = peaks.get("hbm_gbs") or 6650.0
    traffic, traffic_src, ktraffic = None, None, {}
    try:  # DRAM bytes per launch from the committed ncu capture of one step (scripts/ncu_traffic.py writes the file beside the csv)
        tj = json.load(open(os.path.join(ROOT, "profiles", "r02_kernel_traffic.json")))
        ktraffic = tj.get(args.workload, {})
        g = ktraffic.get("gemm_f16_kernel")
        if g:
            traffic, traffic_src = round(g["dram_bytes_per_launch"]), tj.get("source")
    except Exception:
        pass
    ach = tot_fl / tot_ms * 1e-9 if tot_ms > 0 else 0.0
    gname = "gemm_f16_kernel (tcgen05.mma kind::f16 on fp16 operand planes: every dense projection, forward 3-term, backward single-pass)"
    roof = {"bound": "tensor", "kernel": gname, "achieved": round(ach, 2),
            "peak": peak, "peak_source": "MEASURED_PEAKS.json bf16_tflops_sustained" if "bf16_tflops_sustained" in peaks else "fallback (B200_PROFILING.md)",
            "unit": "TFLOP/s", "frac": round(ach / peak, 4), "traffic": traffic, "traffic_unit": "DRAM bytes per launch (ncu)", "traffic_source": traffic_src,
            "hw_tflops": round(hw_fl / tot_ms * 1e-9, 2) if tot_ms > 0 else None, "hw_frac": round(hw_fl / tot_ms * 1e-9 / peak, 4) if tot_ms > 0 else None,
            "algorithmic_bytes_per_launch": round(sum(t[3] for t in trace) / max(1, len(trace))), "launches_per_step": len(trace) // 2,
            "gemm_ms_per_step": round(tot_ms / 2, 3), "step_ms_instrumented": round(step_ms, 3),
            # what the large products are bound by: the L2 -> shared-memory fill of the TMA producers; the chip-wide
            # L2 throughput cap of the microarchitecture notes is ~6300 B/cycle at 1.095 GHz for TMA and LDG alike
            "smem_fill": {"bytes_per_launch": round(sum((t[5] if len(t) > 5 else 0.0) for t in trace) / max(1, len(trace))),
                          "achieved_gbs": round(sum((t[5] if len(t) > 5 else 0.0) for t in trace) / tot_ms * 1e-6, 1) if tot_ms > 0 else None,
                          "l2_throughput_cap_gbs": 6900.0,
                          "what": "bytes staged into shared memory by the GEMM's TMA producers / event-timed GEMM time (all launches, the launch-bound small ones included)"},
            "note": "achieved = ALGORITHMIC flops (2MNK per product) / event-timed launch duration; hw_tflops counts the three tensor-core passes of "
                    "the 3-term forward products (hi.hi + lo.hi + hi.lo: what keeps the gradients within 2e-3 of fp64), i.e. what the tensor "
                    "pipe actually executed; launch durations from two extra eager steps with the side streams off, events on the launching stream"}
    # the other kernels of the step, each against the measured HBM copy bandwidth
    agg = {}
    for e0, e1, name, nb in ktrace:
        d = agg.setdefault(name.replace("b200iq_", ""), [0, 0.0, 0.0])
        d[0] += 1
        d[1] += e0.elapsed_time(e1)
        d[2] += nb
    all_ms = tot_ms + sum(v[1] for v in agg.values())   # event-timed launches of the two steps (GEMMs + every traced kernel)
    kernels = [{"kernel": "gemm_f16", "bound": "tensor", "launches_per_step": len(trace) // 2, "ms_per_step": round(tot_ms / 2, 3),
                "achieved": round(ach, 1), "unit": "TFLOP/s", "peak": peak, "frac": round(ach / peak, 4), "share_of_traced_time": round(tot_ms / all_ms, 3)}]
    for name, (n, ms, nb) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        if ms < 0.004 * all_ms:
            continue
        gbs = nb / ms * 1e-6
        kt = ktraffic.get(name)
        kernels.append({"kernel": name, "bound": "hbm", "launches_per_step": n // 2, "ms_per_step": round(ms / 2, 3), "achieved": round(gbs, 1),
                        "unit": "GB/s", "peak": hbm, "frac": round(gbs / hbm, 4), "share_of_traced_time": round(ms / all_ms, 3),
                        "algorithmic_bytes_per_launch": round(nb / n), "traffic": round(kt["dram_bytes_per_launch"]) if kt else None})
    roof["kernels"] = kernels
    return roof, detail


def measure(args, cfg, config, dp, rank, world, local_rank, roofline=True, cpu_arm=True):
    """one workload on this rank: device-timed value, end-to-end value, roofline of the dominant kernel, CPU arm"""
    import torch.distributed as dist
    dist_on = world > 1
    from blt_vqg_b200 import ops
    model, ts = build_ours(args, cfg, dp)
    hb = host_batches(cfg, 4, rank)
    db = [{k: v.cuda(non_blocking=True) for k, v in b.items()} for b in hb]
    torch.cuda.synchronize()

    def dev_step(i):
        ts.step(*step_inputs(args.workload, db[i % len(db)]))

    sinks = [torch.zeros(8).pin_memory() for _ in range(2)]
    sink_ev = [torch.cuda.Event() for _ in range(2)]
    seen = []

    def e2e_step(i):
        # every step: H2D of this step's inputs from pinned host memory (inside the call) and a D2H read of its loss terms.
        # The read is double-buffered: step i's terms are consumed on the host after step i+1 has been queued, so the host
        # keeps launching while the device works (an async logger; the reference's .item() blocks instead).
        b = hb[i % len(hb)]
        terms = ts.step(*step_inputs(args.workload, b))
        k = i & 1
        if i >= 2:
            sink_ev[k].synchronize()
            seen.append(float(sinks[k][0]))
        sinks[k].copy_(terms[:8], non_blocking=True)
        sink_ev[k].record()

    def e2e_blocking_step(i):
        b = hb[i % len(hb)]
        terms = ts.step(*step_inputs(args.workload, b))
        sinks[0].copy_(terms[:8], non_blocking=False)    # synchronises like the reference's .item()

    for i in range(max(3, args.warmup)):
        dev_step(i)
    l0 = ops.LAUNCHES
    sampler = ClockSampler(local_rank)
    sampler.start()
    ms = timed_region(dev_step, args.steps, dist_on)
    launches = ops.LAUNCHES - l0
    ms_e2e = timed_region(e2e_step, args.steps, dist_on)
    ms_e2e_blk = timed_region(e2e_blocking_step, args.steps, dist_on)
    # the same end-to-end loop fed through the input pipeline (batch.BatchStager + batch.FeatureCache):
    # int64 token tensors as the reference's collate_fn makes them -> one pinned int32 block -> one copy; the features come
    # from the HBM-resident bf16 cache of the frozen trunk's output (COCO-sized: 123287 images), so a step moves row indices.
    from blt_vqg_b200.batch import BatchStager, FeatureCache
    stager, n_img = BatchStager(), 123287
    cache = FeatureCache(n_img, cfg["FEAT"], dtype=torch.bfloat16)
    cache.rows.copy_(torch.randn(4096, cfg["FEAT"], device="cuda").abs().bfloat16().repeat(n_img // 4096 + 1, 1)[:n_img])
    model.feature_cache = cache
    rb = [{**{k: v.long() for k, v in b.items() if not v.is_floating_point()},
           "rows": torch.randint(0, n_img, (b["feats"].shape[0],), generator=torch.Generator().manual_seed(i))} for i, b in enumerate(hb)]
    staged_bytes = []

    def e2e_cached_step(i):
        b = rb[i % len(rb)]
        sb = stager.stage(b, image_rows=b["rows"]).wait()
        x = step_inputs(args.workload, {"feats": sb.image_rows, "answers": sb.answers, "posteriors": sb.posteriors,
                                        "questions": sb.questions, "alengths": sb.alengths})
        terms = ts.step(*x)
        sb.done()
        staged_bytes.append(sb.h2d_bytes)
        k = i & 1
        if i >= 2:
            sink_ev[k].synchronize()
        sinks[k].copy_(terms[:8], non_blocking=True)
        sink_ev[k].record()

    for i in range(3):
        e2e_cached_step(i)
    ms_e2e_cached = timed_region(e2e_cached_step, args.steps, dist_on)
    del cache, stager
    model.feature_cache = None
    clocks = sampler.stop()
    terms = ts.step(*step_inputs(args.workload, db[0])).cpu()
    finite = bool(torch.isfinite(terms[:4]).all())
    in_sync = None
    if dist_on:  # data-parallel replicas must hold bit-identical weights after the same number of reduced steps
        chk = model.flat.data.double().sum().view(1)
        lo, hi = chk.clone(), chk.clone()
        dist.all_reduce(lo, op=dist.ReduceOp.MIN)
        dist.all_reduce(hi, op=dist.ReduceOp.MAX)
        in_sync = bool((hi - lo).abs().item() == 0.0)

    if rank != 0:
        return None, None
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    roof, detail = (None, None)
    if roofline:
        saved_dp, ts.dp = ts.dp, None  # rank-0-only instrumented pass: no collectives
        roof, detail = kernel_rooflines(ts, model, args, db, peaks)
        ts.dp = saved_dp
    cpu = None
    if cpu_arm and world == 1:
        r = reference_steps(args.workload, cfg, 3, 1, args.cpu_batch)
        cpu = {"value": round(r["value"], 2), "unit": UNIT, "cores": r["cores"], "kind": r["kind"], "ms_per_step": round(r["ms_per_step"], 1),
               "sample": "median of 3 timed steps (after 1 warm-up) of fwd+losses+bwd+clip+Adam at batch %d of the same shapes, fp32 PyTorch on the host "
                         "cores, %s" % (r["batch"], "the reference's own modules from baseline/_ref" if r["kind"] == "reference" else "oracle port (no reference tree reachable)")}
        if not args.no_eager_baseline:
            # the same reference modules as stock PyTorch eager on this B200 at the FULL batch (device-resident inputs)
            eager = {}
            for mode in ("fp32", "tf32", "bf16"):
                try:
                    e = reference_steps(args.workload, cfg, 5, 3, cfg["B"], device="cuda", mode=mode)
                    eager[mode] = {"value": round(e["value"], 1), "ms_per_step": round(e["ms_per_step"], 2), "ms_per_step_min": round(e["ms_per_step_min"], 2)}
                except Exception as ex:  # informational leg: never fail the bench line over it
                    eager[mode] = {"error": "%s: %s" % (type(ex).__name__, str(ex)[:120])}
                torch.cuda.empty_cache()
            torch.backends.cuda.matmul.allow_tf32 = False
            torch.backends.cudnn.allow_tf32 = False
            cpu["eager_b200"] = dict(eager, unit=UNIT, kind=r["kind"],
                                     sample="the reference's modules run by stock PyTorch eager on cuda:0 (cuBLAS / cuDNN / ATen), batch %d, median of 5 timed steps "
                                            "after 3 warm-up; fp32 = TF32 off, tf32 = allow_tf32, bf16 = autocast" % cfg["B"])
    B = cfg["B"]
    h2d = sum(v.numel() * v.element_size() for k, v in hb[0].items() if k in ("feats", "answers", "posteriors", "questions", "alengths"))
    line = {"metric": METRIC, "value": round(B * world * args.steps / (ms * 1e-3), 1), "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": max(3, args.warmup), "ms_per_step": round(ms / args.steps, 3), "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": DTYPES["transformer" if args.workload == "transformer" else "lstm"], "data": "synthetic", "config": config,
            "e2e": {"value": round(B * world * args.steps / (ms_e2e * 1e-3), 1), "unit": UNIT, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": 32,
                    "ms_per_step": round(ms_e2e / args.steps, 3), "loss_read": "every step, double-buffered pinned sink, consumed one step later",
                    "blocking_read_value": round(B * world * args.steps / (ms_e2e_blk * 1e-3), 1),
                    "staged_cached": {"value": round(B * world * args.steps / (ms_e2e_cached * 1e-3), 1), "h2d_bytes_per_step": staged_bytes[-1],
                                      "what": "int64 batch dict -> BatchStager (one pinned int32 block, one copy) + rows of the bf16 FeatureCache (123287 x FEAT in HBM)"}},
            "gpu_launches": launches, "gpu_launches_per_step": launches // max(1, args.steps), "clocks": clocks,
            "cuda_graph": {"enabled": bool(ts.graph), "replays": ts.graph_replays,
                           "what": "the whole step (every stream fork / join, NCCL buckets included) captured once per input shape and replayed; "
                                   "gpu_launches counts the kernel nodes each replay runs"},
            "roofline": roof, "cpu_baseline": cpu, "loss_terms_finite": finite, "replicas_in_sync": in_sync,
            "allreduce_buckets_per_step": (dp.collectives // max(1, ts.t)) if dp is not None else 0,
            "loss_terms": {"rec": round(float(terms[0]), 4), "img": round(float(terms[1]), 4), "kld": round(float(terms[2]), 4), "aux": round(float(terms[3]), 4)}}
    return line, detail


_REAL_STDOUT = None


def emit(line):
    """the ONE JSON line goes to the process's original stdout; everything else (NCCL's version banner, library chatter) was
    re-routed to stderr by main()"""
    data = (json.dumps(line) + "\n").encode()
    if _REAL_STDOUT is None:
        sys.stdout.write(data.decode())
        sys.stdout.flush()
    else:
        os.write(_REAL_STDOUT, data)


def main():
    global _REAL_STDOUT
    args = parse()
    sys.stdout.flush()
    _REAL_STDOUT = os.dup(1)
    os.dup2(2, 1)   # fd 1 -> stderr for the rest of the run (native libraries print banners with printf)
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    cfg, wname = workload_config(args)
    config = {"workload": wname, "per_gpu_batch": cfg["B"], "global_batch": cfg["B"] * world,
              "parallelism": "dp%d" % world, "trunk": "excluded (pooled features in)",
              "l2_policy": "inputs larger than L2: the per-step working set (weights, gradients, Adam moments, activations: > 2 GB) far exceeds the 126 MB L2; 4 rotating input batches"}

    if args.impl == "reference":
        # the reference's own CPU implementation of the path: rank 0 alone, all host threads, the workload's own batch unless the
        # first step predicts that --warmup + --steps would run past the budget (then the batch is cut and reported as run)
        if rank != 0:
            return 0
        steps, warm = max(1, args.steps), max(1, args.warmup)
        r = reference_steps(args.workload, cfg, steps, warm, args.ref_batch or cfg["B"], budget_s=float(os.environ.get("B200IQ_REF_BUDGET_S", "240")))
        config = dict(config, per_gpu_batch=r["batch"], global_batch=r["batch"], parallelism="cpu (rank 0 only)")
        sample = ("fwd+losses+bwd+clip_grad_norm_(5)+Adam, fp32 stock PyTorch on the host cores, %s, batch %d of the workload's shapes, "
                  "%d warm-up + %d timed steps, value from the median step" % (
                      "the reference's own modules (baseline/_ref: models.IQ + TrainIQ.calculate_losses)" if r["kind"] == "reference"
                      else "oracle port (no reference tree reachable)", r["batch"], r["warmup"], r["steps"]))
        line = {"impl": "reference", "metric": METRIC, "value": round(r["value"], 2), "unit": UNIT, "n_gpus": args.gpus, "steps": r["steps"], "warmup": r["warmup"],
                "ms_per_step": round(r["ms_per_step"], 2), "ms_per_step_min": round(r["ms_per_step_min"], 2), "higher_is_better": True, "scaling": "weak",
                "vs_baseline": None, "dtype": "f32", "data": "synthetic", "config": config,
                "cpu_baseline": {"value": round(r["value"], 2), "unit": UNIT, "cores": r["cores"], "kind": r["kind"], "sample": sample,
                                 "min_step_value": round(r["min_step_value"], 2), "max_step_value": round(r["max_step_value"], 2)},
                "e2e": {"value": round(r["value"], 2), "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
        emit(line)
        return 0

    import torch.distributed as dist
    dist_on = world > 1
    torch.cuda.set_device(local_rank)
    dp = None
    if dist_on:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
        from blt_vqg_b200.parallel import GradReducer
        dp = GradReducer()
    line, detail = measure(args, cfg, config, dp, rank, world, local_rank, roofline=not args.no_roofline,
                           cpu_arm=not args.no_cpu_baseline)
    if rank == 0 and args.workload == "transformer" and not args.no_secondary and not args.config:
        # BASELINE.json configs[1] (LSTM decoder, full size) measured the same way, reported beside the headline workload
        import copy
        a2 = copy.copy(args)
        a2.workload, a2.batch, a2.config = "lstm", 0, ""
        cfg2, name2 = workload_config(a2)
        if world == 1:
            l2, _ = measure(a2, cfg2, {"workload": name2}, None, rank, world, local_rank, roofline=not args.no_roofline, cpu_arm=not args.no_cpu_baseline)
            line["other_workloads"] = {"lstm_cfg2": {k: l2[k] for k in ("value", "unit", "ms_per_step", "e2e", "gpu_launches_per_step", "loss_terms", "roofline", "cpu_baseline", "dtype")}}
            line["other_workloads"]["lstm_cfg2"]["note"] = ("the LSTM glue (how the blocks are wired) is not in the reference tree: parity of this variant is pinned at the "
                                                            "block level (DecoderRNN / EncoderRNN / MLP), the wiring follows the call-site contract (SURVEY 8c)")
            line["other_workloads"]["lstm_cfg2"]["workload"] = name2
    if rank != 0:
        if dist_on:
            dist.barrier()
            dist.destroy_process_group()
        return 0
    emit(line)
    if detail:
        sys.stderr.write("GEMM breakdown (top shapes by time, 2 instrumented steps):\n")
        for d in detail:
            sys.stderr.write("  " + json.dumps(d) + "\n")
    if dist_on:
        dist.barrier()
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
