"""GPU diagnostic: gradient error of (a) fp32 CPU oracle, (b) the oracle on the GPU with torch's TF32 matmuls, (c) our exact
path, (d) our TF32 tensor-core path — all against the fp64 CPU oracle."""
import sys, os, warnings
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
warnings.filterwarnings("ignore")
import torch
from blt_vqg_b200 import synth, ops
from blt_vqg_b200.iq import IQ
from oracle import iq_oracle

def oracle_grads(sd, cfg, batch, eps, latent, dtype, device, tf32=False):
    torch.backends.cuda.matmul.allow_tf32 = tf32
    torch.backends.cudnn.allow_tf32 = tf32
    ocfg = dict(num_layers=cfg["NL"], num_heads=cfg["NH"], latent_dim=cfg["L"], image_recon_lambda=0.1, kl_ceiling=0.5, aux_ceiling=1.0, full_kl_step=15000)
    P = {k: (v.to(device=device, dtype=dtype) if v.dtype.is_floating_point else v.to(device)).clone() for k, v in sd.items()}
    for k, v in P.items():
        if v.dtype.is_floating_point and "running" not in k: v.requires_grad_(True)
    bn = (P["encoder_cnn.bn.running_mean"].clone(), P["encoder_cnn.bn.running_var"].clone())
    b = {k: v.to(device) for k, v in batch.items()}
    out, zl, kld, img, recon = iq_oracle.iq_forward(P, ocfg, b["feats"].to(dtype), b["answers"], b["posteriors"], b["questions"], eps=eps.to(device=device, dtype=dtype), latent_mode=latent, training=True, bn_state=bn)
    loss, tr = iq_oracle.iq_losses(ocfg, out, zl, kld, img, recon, b["questions"], 7000, latent)
    loss.backward()
    return {k: v.grad.detach().double().cpu() for k, v in P.items() if v.requires_grad and v.grad is not None}, float(loss)

def report(tag, g, truth):
    errs = []; num = den = 0.0
    for n, gt in truth.items():
        if n not in g or n == "encoder_cnn.cnn.fc.bias": continue
        d = (g[n] - gt).norm().item(); r = gt.norm().item()
        num += d * d; den += r * r
        errs.append((d / (r + 1e-30), n))
    errs.sort(reverse=True)
    print("   %-22s global %.2e | worst %s" % (tag, (num / den) ** 0.5, ["%s %.1e" % (n.replace("answer_encoder.", "ae.").replace("multi_head_attention", "mha").replace("positionwise_feed_forward", "ffn")[-34:], e) for e, n in errs[:3]]), flush=True)

def run(name, B=None, latent=True):
    cfg = dict(synth.CONFIGS[name])
    if B: cfg["B"] = B
    batch = synth.make_batch(cfg)
    torch.manual_seed(3)
    m0 = IQ(latent, synth.SynthVocab(cfg["V"]), synth.args_for(cfg))
    sd = {k: v.clone() for k, v in m0.state_dict().items()}
    names = set(m0.arena.names())
    eps = torch.randn(cfg["B"], cfg["L"])
    truth, l64 = oracle_grads(sd, cfg, batch, eps, latent, torch.float64, "cpu")
    truth = {k: v for k, v in truth.items() if k in names}
    print("%s B=%d latent=%d loss64=%.6f" % (name, cfg["B"], latent, l64))
    report("oracle fp32 cpu", oracle_grads(sd, cfg, batch, eps, latent, torch.float32, "cpu")[0], truth)
    report("oracle fp32 gpu", oracle_grads(sd, cfg, batch, eps, latent, torch.float32, "cuda")[0], truth)
    report("oracle torch-tf32 gpu", oracle_grads(sd, cfg, batch, eps, latent, torch.float32, "cuda", tf32=True)[0], truth)
    for simt in (True, False):
        m = IQ(latent, synth.SynthVocab(cfg["V"]), synth.args_for(cfg)).cuda().train()
        m.load_state_dict(sd); m.attn_dropout = m.relu_dropout = 0.0
        with ops.gemm_mode(simt=simt):
            t = m.forward_losses(batch["feats"], batch["answers"], batch["posteriors"], batch["questions"], kliter=7000, eps=eps.cuda()).cpu()
        g = {n: m.arena.view_of(m.unscaled_grads(), n).double().cpu() for n in m.arena.names()}
        report("ours exact" if simt else "ours tf32", g, truth)

run("tiny"); run("small"); run("cfg1", B=32); run("cfg1", B=32, latent=False)
