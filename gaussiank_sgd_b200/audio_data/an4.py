#!/usr/bin/env python
"""Prepare the AN4 corpus (the reference's audio_data/an4.py:1-87): unpack ``an4_raw.bigendian.tar.gz``,
convert the big-endian 16 kHz raw files to wav, write the transcripts, build the two manifests.
There is no network here, so the tarball must already be on disk (``--tar-path``); the conversion is done in
numpy instead of ``sox``."""
import argparse
import os
import tarfile
import wave

import numpy as np

from .utils import create_manifest


def raw_to_wav(raw_path, wav_path, rate=16000):
    pcm = np.fromfile(raw_path, dtype='>i2').astype('<i2')
    with wave.open(wav_path, 'wb') as w:
        w.setnchannels(1); w.setsampwidth(2); w.setframerate(rate)
        w.writeframes(pcm.tobytes())


def format_split(root, target_dir, tag, wav_folder):
    txt_dir = os.path.join(target_dir, tag, 'an4', 'txt')
    wav_dir = os.path.join(target_dir, tag, 'an4', 'wav')
    os.makedirs(txt_dir, exist_ok=True)
    os.makedirs(wav_dir, exist_ok=True)
    with open(os.path.join(root, 'etc', 'an4_%s.fileids' % tag)) as f:
        ids = [l.strip() for l in f if l.strip()]
    with open(os.path.join(root, 'etc', 'an4_%s.transcription' % tag)) as f:
        trans = [l.strip() for l in f if l.strip()]
    for fid, line in zip(ids, trans):
        text = line.split('(')[0].strip('<s>').split('<')[0].strip().upper()
        name = os.path.basename(fid)
        raw_to_wav(os.path.join(root, 'wav', fid + '.raw'), os.path.join(wav_dir, name + '.wav'))
        with open(os.path.join(txt_dir, name + '.txt'), 'w') as t:
            t.write(text)
    return os.path.join(target_dir, tag)


def main():
    ap = argparse.ArgumentParser(description='Processes an4.')
    ap.add_argument('--tar-path', required=True, help='an4_raw.bigendian.tar.gz (download it where there is network)')
    ap.add_argument('--target-dir', default='an4_dataset/')
    ap.add_argument('--min-duration', default=1, type=int)
    ap.add_argument('--max-duration', default=15, type=int)
    a = ap.parse_args()
    with tarfile.open(a.tar_path) as tar:
        tar.extractall(a.target_dir)
    root = os.path.join(a.target_dir, 'an4')
    train = format_split(root, a.target_dir, 'train', 'an4_clstk')
    test = format_split(root, a.target_dir, 'test', 'an4test_clstk')
    print(create_manifest(train, 'an4_train_manifest.csv', a.min_duration, a.max_duration), 'training utterances')
    print(create_manifest(test, 'an4_val_manifest.csv'), 'validation utterances')


if __name__ == '__main__':
    main()
