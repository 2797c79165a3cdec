"""Synthetic datasets with the shapes of the reference's workloads.

There is no network on the build or GPU boxes, so every dataset the reference downloads or reads from
site-specific paths (dl_trainer.py:295-483) has a deterministic synthetic stand-in of the same shape and
dtype.  Samples are generated from a per-index seed, so every rank sees the same "dataset" and a
``DistributedSampler`` shards it exactly like real data.  Labels are a deterministic function of the
sample, which makes the tasks learnable (used by the convergence smoke tests).
"""
from __future__ import annotations

import torch
from torch.utils.data import Dataset

SIZES = {  # (train, test) sample counts of the real datasets
    "cifar10": (50000, 10000), "mnist": (60000, 10000), "imagenet": (1281167, 50000),
    "ptb": (929589, 73760), "an4": (897, 130),
}


class SyntheticImageDataset(Dataset):
    """Gaussian images whose class shifts the mean of one channel-quadrant (learnable)."""

    def __init__(self, n: int, shape, num_classes: int, seed: int = 0, pool: int = 512):
        self.n, self.shape, self.num_classes = n, tuple(shape), num_classes
        pool = min(pool, n)
        g = torch.Generator().manual_seed(seed)
        self.labels = torch.randint(0, num_classes, (pool,), generator=g)
        self.images = torch.randn((pool,) + self.shape, generator=g)
        c, h, w = self.shape
        for i in range(pool):
            y = int(self.labels[i])
            ch, q = y % c, (y // c) % 4
            hs = slice(0, h // 2) if q < 2 else slice(h // 2, h)
            ws = slice(0, w // 2) if q % 2 == 0 else slice(w // 2, w)
            self.images[i, ch, hs, ws] += 1.5 + 0.25 * (y // (4 * c))
        self.pool = pool

    def __len__(self):
        return self.n

    def __getitem__(self, i):
        j = i % self.pool
        return self.images[j], int(self.labels[j])


class SyntheticTokenDataset(Dataset):
    """PTB-shaped token stream cut into (num_steps,) windows with next-token targets, like
    ptb_reader.TrainDataset (ptb_reader.py:56-78).  Tokens follow a fixed random bigram chain so a
    language model can reduce its perplexity."""

    def __init__(self, n_tokens: int, vocab: int, batch_size: int, num_steps: int, seed: int = 0):
        g = torch.Generator().manual_seed(seed)
        nxt = torch.randint(0, vocab, (vocab, 4), generator=g)
        choice = torch.randint(0, 4, (n_tokens,), generator=g)
        toks = torch.empty(n_tokens, dtype=torch.long)
        t = 0
        raw = toks.numpy()
        nx, chn = nxt.numpy(), choice.numpy()
        for i in range(n_tokens):
            raw[i] = t
            t = nx[t, chn[i]]
        self.raw = toks
        self.num_steps = num_steps
        sample_len = (n_tokens - 1) // num_steps
        self.len = sample_len - sample_len % batch_size

    def __len__(self):
        return self.len

    def __getitem__(self, i):
        s = i * self.num_steps
        return self.raw[s:s + self.num_steps], self.raw[s + 1:s + self.num_steps + 1]


class SyntheticSpectrogramDataset(Dataset):
    """AN4-shaped utterances: (1, 161, T) log-spectrograms with T in [lo, hi] and 29-symbol
    transcripts (audio_data interface reconstructed in SURVEY.md §2.1)."""

    def __init__(self, n: int, n_labels: int = 29, freq: int = 161, t_lo: int = 100, t_hi: int = 400, seed: int = 0):
        g = torch.Generator().manual_seed(seed)
        self.n, self.freq, self.n_labels = n, freq, n_labels
        self.lengths = torch.randint(t_lo, t_hi + 1, (n,), generator=g)
        self.seed = seed

    def __len__(self):
        return self.n

    def __getitem__(self, i):
        T = int(self.lengths[i])
        g = torch.Generator().manual_seed(self.seed * 1000003 + i)
        spect = torch.randn(self.freq, T, generator=g)
        n_tok = max(1, T // 20)
        transcript = torch.randint(1, self.n_labels, (n_tok,), generator=g).int()
        # imprint the transcript so that it is recoverable from the spectrogram
        for j, tok in enumerate(transcript.tolist()):
            spect[(tok * 5) % self.freq, j * 20:(j + 1) * 20] += 3.0
        return spect, transcript


class PinnedBatchPool:
    """Pre-collated mini-batches of a synthetic image dataset in pinned host memory, cycled.

    A ``DataLoader`` collating 32 x 3 x 224 x 224 floats on the main thread costs more host time per
    batch than a B200 needs for the whole training step; with synthetic data there is nothing to decode,
    so the batches are built once.  Every step still performs its own host->device copy from pinned
    memory.  Rank r of P takes samples r, r+P, ... like a ``DistributedSampler`` without shuffling."""

    def __init__(self, dataset: SyntheticImageDataset, batch_size: int, rank: int = 0, nworkers: int = 1,
                 nbatches: int = 8, pin: bool = True):
        self.dataset, self.batch_size = dataset, batch_size
        self.len = max(len(dataset) // (batch_size * nworkers), 1)
        nbatches = max(1, min(nbatches, self.len))
        self.batches = []
        for b in range(nbatches):
            idx = (torch.arange(batch_size) * nworkers + rank + b * batch_size * nworkers) % dataset.pool
            x, y = dataset.images[idx].contiguous(), dataset.labels[idx].contiguous()
            if pin and torch.cuda.is_available():
                x, y = x.pin_memory(), y.pin_memory()
            self.batches.append((x, y))

    def __len__(self):
        return self.len

    def __iter__(self):
        for i in range(self.len):
            yield self.batches[i % len(self.batches)]
