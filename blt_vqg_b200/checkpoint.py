"""Checkpoints in the reference's formats <-> the flat arenas.

The reference writes two kinds of file:
  * Lightning `.ckpt` from `trainer.save_checkpoint` (`CheckpointEveryNSteps`, train_iq.py:275-309,373): a dict whose
    `"state_dict"` holds the `TrainIQ` module, i.e. every `IQ` tensor under the prefix `model.` (train_iq.py:53), shared
    sub-modules under each of their attribute paths (`embedding`, `answer_encoder.embedding`, `decoder.embedding`, ...;
    models/iq.py:32-43), and the frozen ResNet trunk under `model.encoder_cnn.cnn.*`;
  * the legacy `torch.save(model.state_dict())` (old_train_iq.py:461-469): the same keys without the prefix.
Both load into `IQ` / `LSTMIQ` here, and `save_checkpoint` writes a file the reference's `load_from_checkpoint` /
`load_state_dict` accepts for everything downstream of the trunk.  The reference wires no resume path, so the
optimizer state travels in this framework's own layout (the flat Adam arenas) under the key `b200iq_trainer`; a Lightning
`optimizer_states` entry written by the reference is ignored (its per-parameter indices count the trunk's tensors).
"""
import torch

PREFIX = "model."          # TrainIQ.model (train_iq.py:53)
TRUNK = "encoder_cnn.cnn."  # frozen torchvision ResNet (models/encoder_cnn.py:17-19); its replaced `fc` is the trained head
VERSION_KEY = "pytorch-lightning_version"


def model_state_dict(ckpt):
    """the `IQ` state_dict inside a Lightning checkpoint, a legacy state_dict file, or an already-plain state_dict"""
    sd = ckpt.get("state_dict", ckpt) if isinstance(ckpt, dict) else ckpt
    if any(k.startswith(PREFIX) for k in sd):
        sd = {k[len(PREFIX):]: v for k, v in sd.items() if k.startswith(PREFIX)}
    return sd


def split_trunk(sd):
    """(head state_dict, trunk state_dict): `encoder_cnn.cnn.fc.*` is the trained projection and stays with the head"""
    head, trunk = {}, {}
    for k, v in sd.items():
        if k.startswith(TRUNK) and not k.startswith(TRUNK + "fc."):
            trunk[k[len(TRUNK):]] = v
        else:
            head[k] = v
    return head, trunk


def load_checkpoint(ckpt, model, train_step=None, strict=True, map_location="cpu"):
    """`ckpt`: path or loaded dict.  Returns (missing, unexpected) like `load_state_dict`.  Trunk tensors go to the attached
    trunk (`IQ.attach_trunk`) when there is one and are dropped otherwise (the trunk is outside the hot path)."""
    if isinstance(ckpt, (str, bytes)) or hasattr(ckpt, "__fspath__"):
        ckpt = torch.load(ckpt, map_location=map_location, weights_only=False)
    head, trunk = split_trunk(model_state_dict(ckpt))
    missing, unexpected = model.load_state_dict(head, strict=strict)
    t = getattr(model, "trunk", None)
    if trunk and t is not None and hasattr(t, "load_state_dict"):
        t.load_state_dict(trunk, strict=False)
    extra = ckpt.get("b200iq_trainer") if isinstance(ckpt, dict) else None
    if train_step is not None:
        if extra is not None:
            train_step.exp_avg.copy_(extra["exp_avg"])
            train_step.exp_avg_sq.copy_(extra["exp_avg_sq"])
            train_step.t, train_step.iter, train_step.kliter = int(extra["t"]), int(extra["iter"]), int(extra["kliter"])
            model.step_idx = int(extra.get("step_idx", train_step.iter))
        elif isinstance(ckpt, dict) and "global_step" in ckpt:
            # a checkpoint written by the reference: only the global step is known (TrainIQ.iter / kliter are not saved and the
            # file carries no mode flag).  The LR schedule needs iter; the KL annealing and the latent switch follow from it
            # (train_iq.py:108-111: the switch happens at iter == num_pretraining_steps, kliter counts from there)
            train_step.iter = int(ckpt["global_step"])
            pre = getattr(getattr(model, "args", None), "num_pretraining_steps", -1)
            if pre is not None and pre >= 0 and train_step.iter >= pre:
                train_step.kliter = train_step.iter - pre
                if hasattr(model, "switch_GVT_train_mode"):
                    model.switch_GVT_train_mode(True)
            train_step._seed_ctr = None
            model.step_idx = train_step.iter
        if getattr(train_step, "_seed_ctr", None) is not None:
            train_step._seed_ctr.fill_(int(model.step_idx))   # the device-resident pass counter of the CUDA-graph step
    if isinstance(ckpt, dict) and "b200iq_latent_transformer" in ckpt:
        model.switch_GVT_train_mode(bool(ckpt["b200iq_latent_transformer"]))
    return missing, unexpected


def sync_batchnorm_stats(model, group=None):
    """average the BatchNorm running statistics over the data-parallel ranks (they are per-rank, like DDP's — where
    broadcast_buffers=True then makes rank 0's win; the mean is the better estimate): call before rank 0 writes a checkpoint"""
    import torch.distributed as dist
    if dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1:
        for name in ("bn_running_mean", "bn_running_var"):
            t = getattr(model, name, None)
            if t is not None:
                dist.all_reduce(t, group=group)
                t.div_(dist.get_world_size(group))


def make_checkpoint(model, train_step=None, epoch=0, hyper_parameters=None):
    """a Lightning-shaped checkpoint dict of `model` (the keys `TrainIQ.load_from_checkpoint` / `IQ.load_state_dict` read).
    Under data parallelism every rank calls it (the BatchNorm statistics are averaged over the ranks first)."""
    sync_batchnorm_stats(model)
    sd = {PREFIX + k: v.detach().cpu().clone() for k, v in model.state_dict().items()}
    t = getattr(model, "trunk", None)
    if t is not None and hasattr(t, "state_dict"):
        for k, v in t.state_dict().items():
            if not k.startswith("fc."):
                sd[PREFIX + TRUNK + k] = v.detach().cpu().clone()
    ckpt = {"epoch": int(epoch), "global_step": int(train_step.iter) if train_step is not None else 0, VERSION_KEY: "1.1.0",
            "state_dict": sd, "b200iq_latent_transformer": bool(getattr(model, "latent_transformer", False))}
    if hyper_parameters is not None:
        ckpt["hyper_parameters"] = dict(hyper_parameters)
    if train_step is not None:
        ckpt["b200iq_trainer"] = {"exp_avg": train_step.exp_avg.detach().cpu().clone(),
                                  "exp_avg_sq": train_step.exp_avg_sq.detach().cpu().clone(),
                                  "t": train_step.t, "iter": train_step.iter, "kliter": train_step.kliter,
                                  "step_idx": int(getattr(model, "step_idx", 0))}
    return ckpt


def save_checkpoint(path, model, train_step=None, epoch=0, hyper_parameters=None):
    torch.save(make_checkpoint(model, train_step, epoch, hyper_parameters), path)
