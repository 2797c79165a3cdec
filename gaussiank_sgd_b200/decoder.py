"""CTC decoders and error rates for the AN4 workload.

The reference imports ``decoder.GreedyDecoder`` (dl_trainer.py:457, 482-483, 787-797) but does not ship the
module; the interface is reconstructed from its use: ``decode(probs[N,T,C], sizes) -> (strings, offsets)``,
``convert_to_strings(list_of_int_sequences)``, ``wer(a, b)`` / ``cer(a, b)`` as edit distances."""
from __future__ import annotations

from typing import List, Sequence

import torch


def _edit_distance(a: Sequence, b: Sequence) -> int:
    """Levenshtein distance with two rolling rows."""
    if len(a) < len(b):
        a, b = b, a
    prev = list(range(len(b) + 1))
    for i, x in enumerate(a, 1):
        cur = [i]
        for j, y in enumerate(b, 1):
            cur.append(min(prev[j] + 1, cur[j - 1] + 1, prev[j - 1] + (x != y)))
        prev = cur
    return prev[-1]


class Decoder:
    def __init__(self, labels: str, blank_index: int = 0):
        self.labels = labels
        self.int_to_char = dict(enumerate(labels))
        self.blank_index = blank_index
        self.space_index = labels.index(' ') if ' ' in labels else len(labels)

    def wer(self, s1: str, s2: str) -> int:
        """Word-level edit distance between two transcripts."""
        return _edit_distance(s1.split(), s2.split())

    def cer(self, s1: str, s2: str) -> int:
        """Character-level edit distance (spaces removed)."""
        return _edit_distance(s1.replace(' ', ''), s2.replace(' ', ''))

    def decode(self, probs, sizes=None):
        raise NotImplementedError


class GreedyDecoder(Decoder):
    def convert_to_strings(self, sequences, sizes=None, remove_repetitions=False, return_offsets=False):
        """Integer sequences -> [[string]] (one-element lists, like a beam of width 1)."""
        strings, offsets = [], []
        for i, seq in enumerate(sequences):
            seq = seq.tolist() if torch.is_tensor(seq) else list(seq)
            n = int(sizes[i]) if sizes is not None else len(seq)
            s, o = self._process(seq[:n], remove_repetitions)
            strings.append([s])
            offsets.append([o])
        return (strings, offsets) if return_offsets else strings

    def _process(self, seq: List[int], remove_repetitions: bool):
        chars, offs = [], []
        for i, c in enumerate(seq):
            if c == self.blank_index:
                continue
            if remove_repetitions and i > 0 and c == seq[i - 1]:
                continue
            chars.append(self.int_to_char[int(c)])
            offs.append(i)
        return ''.join(chars), torch.tensor(offs, dtype=torch.int)

    def decode(self, probs: torch.Tensor, sizes=None):
        """Best path: arg-max per frame, collapse repeats, drop blanks.  ``probs`` is (N, T, C)."""
        best = probs.argmax(dim=2)
        return self.convert_to_strings(best, sizes, remove_repetitions=True, return_offsets=True)
