"""Manifest helpers for the speech corpora (the reference's audio_data/utils.py:1-44)."""
import fnmatch
import os
import wave

import torch.distributed as dist


def wav_duration(path: str) -> float:
    """Seconds of a PCM wav file (the reference shells out to ``soxi -D``)."""
    with wave.open(path, 'rb') as w:
        return w.getnframes() / float(w.getframerate())


def order_and_prune_files(file_paths, min_duration=None, max_duration=None):
    timed = [(p, wav_duration(p)) for p in file_paths]
    if min_duration and max_duration:
        timed = [(p, d) for p, d in timed if min_duration <= d <= max_duration]
    timed.sort(key=lambda t: t[1])
    return [p for p, _ in timed]


def create_manifest(data_path, output_path, min_duration=None, max_duration=None):
    """One ``wav,txt`` line per utterance, shortest first (transcripts live in a parallel txt/ folder)."""
    wavs = [os.path.join(d, f) for d, _dirs, files in os.walk(data_path) for f in fnmatch.filter(files, '*.wav')]
    wavs = order_and_prune_files(wavs, min_duration, max_duration)
    with open(output_path, 'w') as out:
        for w in wavs:
            txt = w.replace('/wav/', '/txt/').replace('.wav', '.txt')
            out.write(os.path.abspath(w) + ',' + os.path.abspath(txt) + '\n')
    return len(wavs)


def reduce_tensor(tensor, world_size):
    rt = tensor.clone()
    dist.all_reduce(rt, op=dist.ReduceOp.SUM)
    rt /= world_size
    return rt
