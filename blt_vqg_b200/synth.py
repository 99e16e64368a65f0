"""Synthetic COCO-VQA-shaped batches for parity tests and bench.py.

Layouts follow utils/data_loader.py of the reference:
  questions (B,T)      = [<start>, w..., <end>, pad...]                      data_loader.py:55-60
  posteriors (B,T+1)   = [<pos>, category, w..., pad...]                      :62-69,84
  answers (B,5)        = [<start>, category, a1(,a2), pad...]                 :71-76,115
  answer_types (B,3)   = [<start>, category, <end>]                           :81
  feats (B,FEAT)       = |N(0,1)|  (post-avgpool trunk features are non-negative)
Reserved ids: utils/train_utils.py:18-37.
"""
from types import SimpleNamespace

import numpy as np
import torch

PAD, SOQ, RESP, EOS, UNK, POS = 0, 1, 2, 3, 4, 5
FIRST_CAT, NUM_CAT = 6, 16
FIRST_WORD = FIRST_CAT + NUM_CAT


class SynthVocab:
    """the attribute surface models/iq.py and decoder_transformer.py read from utils/train_utils.py:Vocabulary"""
    SYM_PAD, SYM_SOQ, SYM_RESP, SYM_EOS, SYM_UNK, SYM_POS = "<pad>", "<start>", "<resp>", "<end>", "<unk>", "<pos>"

    def __init__(self, size):
        words = [self.SYM_PAD, self.SYM_SOQ, self.SYM_RESP, self.SYM_EOS, self.SYM_UNK, self.SYM_POS]
        words += ["w%d" % i for i in range(len(words), size)]
        self.idx2word = dict(enumerate(words))
        self.word2idx = {w: i for i, w in self.idx2word.items()}

    def __len__(self):
        return len(self.word2idx)


def make_args(**kw):
    """argparse namespace of train_iq.py:313-358 restricted to what the model / losses read"""
    d = dict(emb_dim=300, hidden_dim=512, latent_dim=100, pwffn_dim=2048, num_layers=6, num_heads=8, emb_file=None,
             root_dir=".", device="cpu", image_recon_lambda=0.1, kl_ceiling=0.5, aux_ceiling=1.0, full_kl_step=15000,
             num_pretraining_steps=12000, feat_dim=2048, input_mode="ans")
    d.update(kw)
    return SimpleNamespace(**d)


CONFIGS = {
    # tiny: checked-in golden fixtures
    "tiny": dict(B=4, V=200, E=32, H=64, L=16, F=128, NL=2, NH=4, T=8, S=5, FEAT=48),
    "small": dict(B=16, V=1000, E=60, H=128, L=32, F=256, NL=2, NH=4, T=12, S=5, FEAT=256),
    # sizes a real run produces: len(vocab.word2idx) is whatever the frequency threshold left (not a multiple of anything), odd
    # embedding / feature widths — the buffers get padded row pitches, the unaligned GEMM epilogues their generic path
    "odd": dict(B=6, V=203, E=30, H=64, L=16, F=128, NL=2, NH=4, T=9, S=5, FEAT=50),
    # cfg3's widths (head_dim 64, H=512, F=2048, V=15000, T=20) at 2 layers and a small batch: the production kernel variants
    # (CTA-pair GEMM, tensor-core attention, warp-per-row LayerNorm, chunked vocabulary CE) against the host oracle in seconds
    "wide": dict(B=24, V=15000, E=300, H=512, L=100, F=2048, NL=2, NH=8, T=20, S=5, FEAT=2048),
    # cfg5's widths (V = 32000, T = 64, R = 65: the row-block attention kernels, 20 logits chunks) at 2 layers and a small batch
    "long": dict(B=8, V=32000, E=300, H=512, L=100, F=2048, NL=2, NH=8, T=64, S=5, FEAT=2048),
    # BASELINE.json configs[0] sizes on the transformer variant
    "cfg1": dict(B=32, V=5000, E=300, H=512, L=100, F=2048, NL=6, NH=8, T=20, S=5, FEAT=2048),
    # BASELINE.json configs[2]: the single-GPU transformer workload the metric is quoted on
    "cfg3": dict(B=512, V=15000, E=300, H=512, L=100, F=2048, NL=6, NH=8, T=20, S=5, FEAT=2048),
    "cfg5": dict(B=128, V=32000, E=300, H=512, L=100, F=2048, NL=6, NH=8, T=64, S=5, FEAT=2048),
}


# LSTM variant (upstream IQ lineage; SURVEY.md §8d config mapping): H hidden, Z latent, A = att_ff width of the MLPs
LSTM_CONFIGS = {
    "tiny": dict(B=6, V=120, H=32, Z=12, A=40, T=8, S=5, FEAT=48),
    "small": dict(B=24, V=1000, H=128, Z=32, A=96, T=12, S=5, FEAT=256),
    "odd": dict(B=10, V=203, H=32, Z=12, A=40, T=9, S=5, FEAT=50),
    "cfg1": dict(B=32, V=5000, H=512, Z=100, A=512, T=20, S=5, FEAT=2048),     # BASELINE.json configs[0]
    "cfg2": dict(B=256, V=15000, H=300, Z=100, A=512, T=20, S=5, FEAT=2048),   # configs[1]: GloVe-300 => H=300
    "cfg4": dict(B=512, V=15000, H=512, Z=100, A=512, T=20, S=5, FEAT=2048),   # configs[3] per-GPU shard
}


def args_for(cfg, **kw):
    return make_args(emb_dim=cfg["E"], hidden_dim=cfg["H"], latent_dim=cfg["L"], pwffn_dim=cfg["F"],
                     num_layers=cfg["NL"], num_heads=cfg["NH"], feat_dim=cfg["FEAT"], **kw)


def make_batch(cfg, seed=1234, rank=0, B=None):
    """dict of CPU tensors (int64 ids like the reference's collate_fn, data_loader.py:168-172)"""
    B = B or cfg["B"]
    V, T, S, FEAT = cfg["V"], cfg["T"], cfg["S"], cfg["FEAT"]
    rs = np.random.RandomState(seed + rank)
    feats = np.abs(rs.randn(B, FEAT)).astype(np.float32)
    mean_words = 6 if T <= 20 else 24
    qlen = np.clip(3 + rs.poisson(mean_words, B), 4, T)
    # Zipf(1.1)-distributed word ids truncated to [FIRST_WORD, V)
    nw = V - FIRST_WORD
    ranks = np.arange(1, nw + 1, dtype=np.float64)
    pz = ranks ** -1.1
    pz /= pz.sum()
    questions = np.zeros((B, T), np.int64)
    posteriors = np.zeros((B, T + 1), np.int64)
    answers = np.zeros((B, S), np.int64)
    atypes = np.zeros((B, 3), np.int64)
    cats = FIRST_CAT + rs.randint(0, NUM_CAT, B)
    for i in range(B):
        n = int(qlen[i]) - 2
        w = FIRST_WORD + rs.choice(nw, n, p=pz)
        questions[i, 0] = SOQ
        questions[i, 1:1 + n] = w
        questions[i, 1 + n] = EOS
        posteriors[i, 0] = POS
        posteriors[i, 1] = cats[i]
        posteriors[i, 2:2 + n] = w
        na = 1 + rs.randint(0, 2)
        answers[i, 0] = SOQ
        answers[i, 1] = cats[i]
        answers[i, 2:2 + na] = FIRST_WORD + rs.choice(nw, na, p=pz)
        atypes[i] = (SOQ, cats[i], EOS)
    return {
        "feats": torch.from_numpy(feats), "questions": torch.from_numpy(questions),
        "posteriors": torch.from_numpy(posteriors), "answers": torch.from_numpy(answers),
        "answer_types_for_input": torch.from_numpy(atypes), "qlen": torch.from_numpy(qlen),
        "alengths": torch.from_numpy((answers != PAD).sum(1)),  # utils/train_utils.py:127-141 process_lengths
    }
