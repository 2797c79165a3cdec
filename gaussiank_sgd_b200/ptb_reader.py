"""Penn Treebank reader (the reference's ptb_reader.py:1-102): frequency-ranked vocabulary built from the
training file, ``<eos>`` at line ends, and datasets that cut the id stream into ``num_steps`` windows with
next-token targets."""
from __future__ import annotations

import collections
import os
from typing import Dict, List, Tuple

import numpy as np
from torch.utils.data import Dataset

EOS = "<eos>"


def read_words(filename: str) -> List[str]:
    with open(filename, "r") as f:
        return f.read().replace("\n", EOS).split()


def build_vocab(filename: str) -> Tuple[Dict[str, int], Dict[int, str]]:
    """Ids by descending frequency, ties alphabetically (ptb_reader.py:14-24)."""
    counts = collections.Counter(read_words(filename))
    ranked = sorted(counts, key=lambda w: (-counts[w], w))
    word_to_id = {w: i for i, w in enumerate(ranked)}
    return word_to_id, {i: w for w, i in word_to_id.items()}


def file_to_word_ids(filename: str, word_to_id: Dict[str, int]) -> List[int]:
    return [word_to_id[w] for w in read_words(filename) if w in word_to_id]


def ptb_raw_data(data_path: str = None, prefix: str = "ptb"):
    """(train_ids, valid_ids, test_ids, word_to_id, id_to_word) from ``<prefix>.{train,valid,test}.txt``."""
    paths = [os.path.join(data_path, "%s.%s.txt" % (prefix, split)) for split in ("train", "valid", "test")]
    word_to_id, id_to_word = build_vocab(paths[0])
    train, valid, test = (file_to_word_ids(p, word_to_id) for p in paths)
    return train, valid, test, word_to_id, id_to_word


class _WindowDataset(Dataset):
    """Item i = (ids[i*T:(i+1)*T], ids[i*T+1:(i+1)*T+1]); the length is truncated to a multiple of the batch
    size so that every batch is full (ptb_reader.py:56-102)."""

    def __init__(self, raw_data, batch_size: int, num_steps: int):
        self.raw_data = np.asarray(raw_data, dtype=np.int64)
        self.num_steps, self.batch_size = num_steps, batch_size
        windows = (len(self.raw_data) - 1) // num_steps
        self.sample_len = windows - windows % batch_size

    def __len__(self):
        return self.sample_len

    def __getitem__(self, idx):
        s = idx * self.num_steps
        return self.raw_data[s:s + self.num_steps], self.raw_data[s + 1:s + self.num_steps + 1]


class TrainDataset(_WindowDataset):
    pass


class TestDataset(_WindowDataset):
    pass
