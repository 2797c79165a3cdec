"""Speech data pipeline for AN4 / LibriSpeech manifests.

The reference imports this module (audio_data/__init__.py:1, dl_trainer.py:456-481) but the file is not in
its tree; it is rebuilt here from the way it is used: ``SpectrogramDataset(audio_conf, manifest_filepath,
labels, normalize, augment)``, ``AudioDataLoader`` whose batches are ``(inputs[N,1,F,T], targets (1-D,
concatenated), input_percentages[N], target_sizes[N])`` sorted by length, ``BucketingSampler`` /
``DistributedBucketingSampler`` that batch utterances of similar length.  Audio is read with ``scipy.io.wavfile``
and turned into log-magnitude STFT spectrograms with ``torch.stft`` (F = 161 at 16 kHz / 20 ms windows)."""
from __future__ import annotations

import math
from typing import List

import numpy as np
import torch
from torch.utils.data import DataLoader, Dataset, Sampler

WINDOWS = {'hamming': torch.hamming_window, 'hann': torch.hann_window, 'blackman': torch.blackman_window,
           'bartlett': torch.bartlett_window}


def load_audio(path: str) -> torch.Tensor:
    from scipy.io import wavfile
    _rate, sound = wavfile.read(path)
    sound = sound.astype('float32') / 32767.0
    if sound.ndim > 1:
        sound = sound[:, 0] if sound.shape[1] == 1 else sound.mean(axis=1)
    return torch.from_numpy(np.ascontiguousarray(sound))


class SpectrogramParser:
    def __init__(self, audio_conf: dict, normalize: bool = False, augment: bool = False):
        self.sample_rate = audio_conf['sample_rate']
        self.window_size = audio_conf['window_size']
        self.window_stride = audio_conf['window_stride']
        self.window = WINDOWS.get(audio_conf.get('window', 'hamming'), torch.hamming_window)
        self.normalize, self.augment = normalize, augment
        self.noise_prob = audio_conf.get('noise_prob', 0.0) if audio_conf.get('noise_dir') else 0.0
        self.noise_levels = audio_conf.get('noise_levels', (0.0, 0.5))

    def spectrogram(self, y: torch.Tensor) -> torch.Tensor:
        n_fft = int(self.sample_rate * self.window_size)
        hop = int(self.sample_rate * self.window_stride)
        if self.augment:                       # light tempo / gain perturbation
            gain = 10 ** (float(np.random.uniform(-6, 8)) / 20)
            y = y * gain
        d = torch.stft(y, n_fft=n_fft, hop_length=hop, win_length=n_fft, window=self.window(n_fft),
                       return_complex=True)
        spect = torch.log1p(d.abs())
        if self.normalize:
            spect = (spect - spect.mean()) / (spect.std() + 1e-8)
        return spect

    def parse_audio(self, audio_path: str) -> torch.Tensor:
        return self.spectrogram(load_audio(audio_path))

    def parse_transcript(self, transcript_path: str):
        raise NotImplementedError


class SpectrogramDataset(Dataset, SpectrogramParser):
    """Manifest lines are ``/path/to/audio.wav,/path/to/transcript.txt``."""

    def __init__(self, audio_conf, manifest_filepath, labels, normalize=False, augment=False):
        with open(manifest_filepath) as f:
            self.ids = [line.strip().split(',') for line in f if line.strip()]
        self.size = len(self.ids)
        self.labels_map = {c: i for i, c in enumerate(labels)}
        SpectrogramParser.__init__(self, audio_conf, normalize, augment)

    def __getitem__(self, index):
        audio_path, transcript_path = self.ids[index]
        return self.parse_audio(audio_path), self.parse_transcript(transcript_path)

    def parse_transcript(self, transcript_path):
        with open(transcript_path, 'r', encoding='utf8') as f:
            transcript = f.read().replace('\n', '')
        return torch.tensor([self.labels_map[c] for c in transcript if c in self.labels_map], dtype=torch.int)

    def __len__(self):
        return self.size


class SyntheticSpeechDataset(Dataset):
    """AN4-shaped stand-in (161 x T log-spectrograms, 29-symbol transcripts) used when the corpus is absent."""

    def __init__(self, n: int, labels: str, seed: int = 0, t_lo: int = 100, t_hi: int = 400):
        from ..data.synthetic import SyntheticSpectrogramDataset
        self.inner = SyntheticSpectrogramDataset(n, n_labels=len(labels), seed=seed, t_lo=t_lo, t_hi=t_hi)
        self.size = n
        self.lengths = self.inner.lengths

    def __len__(self):
        return self.size

    def __getitem__(self, i):
        return self.inner[i]


def _collate_fn(batch):
    """Sort by length (longest first), zero-pad to the longest utterance."""
    batch = sorted(batch, key=lambda s: s[0].size(1), reverse=True)
    freq, max_t = batch[0][0].size(0), batch[0][0].size(1)
    inputs = torch.zeros(len(batch), 1, freq, max_t)
    percentages = torch.empty(len(batch), dtype=torch.float32)
    sizes = torch.empty(len(batch), dtype=torch.int)
    targets: List[torch.Tensor] = []
    for i, (spect, target) in enumerate(batch):
        t = spect.size(1)
        inputs[i, 0, :, :t] = spect
        percentages[i] = t / float(max_t)
        sizes[i] = len(target)
        targets.append(target.int())
    return inputs, torch.cat(targets), percentages, sizes


class AudioDataLoader(DataLoader):
    def __init__(self, *args, **kwargs):
        super().__init__(*args, **kwargs)
        self.collate_fn = _collate_fn


class BucketingSampler(Sampler):
    """Consecutive (length-ordered) utterances form a batch; the batch order is shuffled per epoch."""

    def __init__(self, data_source, batch_size=1):
        super().__init__()
        self.data_source = data_source
        ids = list(range(len(data_source)))
        lengths = getattr(data_source, 'lengths', None)
        if lengths is not None:
            ids.sort(key=lambda i: int(lengths[i]))
        self.bins = [ids[i:i + batch_size] for i in range(0, len(ids), batch_size)]
        self.bins = [b for b in self.bins if len(b) == batch_size] or self.bins

    def __iter__(self):
        for ids in self.bins:
            yield ids

    def __len__(self):
        return len(self.bins)

    def shuffle(self, epoch):
        g = np.random.RandomState(epoch)
        g.shuffle(self.bins)

    set_epoch = shuffle


class DistributedBucketingSampler(Sampler):
    """Every rank takes every ``num_replicas``-th bucket; the bucket list is padded so that all ranks do the
    same number of iterations."""

    def __init__(self, data_source, batch_size=1, num_replicas=None, rank=None):
        super().__init__()
        self.data_source, self.num_replicas, self.rank = data_source, num_replicas or 1, rank or 0
        base = BucketingSampler(data_source, batch_size)
        self.bins = base.bins
        self.num_samples = int(math.ceil(len(self.bins) / self.num_replicas))
        self.total_size = self.num_samples * self.num_replicas

    def __iter__(self):
        bins = self.bins + self.bins[:self.total_size - len(self.bins)]
        for b in bins[self.rank::self.num_replicas][:self.num_samples]:
            yield b

    def __len__(self):
        return self.num_samples

    def shuffle(self, epoch):
        g = np.random.RandomState(epoch)        # same permutation on every rank
        perm = g.permutation(len(self.bins))
        self.bins = [self.bins[i] for i in perm]

    set_epoch = shuffle
