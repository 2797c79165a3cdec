#!/usr/bin/env python
"""Prepare LibriSpeech subsets (the reference's audio_data/librispeech.py:1-113) from tarballs that are already
on disk: flac -> 16 kHz wav, one upper-case transcript file per utterance, then the manifests.  Decoding
flac needs the ``soundfile`` package or a ``sox`` binary; neither ships with this image, so the script checks
and says so."""
import argparse
import os
import shutil
import subprocess
import tarfile

from .utils import create_manifest


def flac_to_wav(src, dst, rate=16000):
    try:
        import soundfile as sf
        data, sr = sf.read(src)
        sf.write(dst, data, sr)
        return
    except ImportError:
        pass
    if shutil.which('sox') is None:
        raise RuntimeError('need the soundfile package or sox to decode flac')
    subprocess.check_call(['sox', src, '-r', str(rate), '-b', '16', '-c', '1', dst])


def process_split(extracted, out_dir):
    wav_dir, txt_dir = os.path.join(out_dir, 'wav'), os.path.join(out_dir, 'txt')
    os.makedirs(wav_dir, exist_ok=True)
    os.makedirs(txt_dir, exist_ok=True)
    for d, _dirs, files in os.walk(extracted):
        for f in files:
            if f.endswith('.trans.txt'):
                with open(os.path.join(d, f)) as t:
                    for line in t:
                        uid, text = line.strip().split(' ', 1)
                        flac_to_wav(os.path.join(d, uid + '.flac'), os.path.join(wav_dir, uid + '.wav'))
                        with open(os.path.join(txt_dir, uid + '.txt'), 'w') as o:
                            o.write(text.upper())


def main():
    ap = argparse.ArgumentParser(description='Processes LibriSpeech tarballs.')
    ap.add_argument('--tars', nargs='+', required=True, help='e.g. train-clean-100.tar.gz dev-clean.tar.gz')
    ap.add_argument('--target-dir', default='LibriSpeech_dataset/')
    ap.add_argument('--min-duration', default=1, type=int)
    ap.add_argument('--max-duration', default=15, type=int)
    a = ap.parse_args()
    for tar_path in a.tars:
        name = os.path.basename(tar_path).split('.tar')[0]
        tmp = os.path.join(a.target_dir, '_extract', name)
        with tarfile.open(tar_path) as tar:
            tar.extractall(tmp)
        out = os.path.join(a.target_dir, name)
        process_split(tmp, out)
        is_train = name.startswith('train')
        n = create_manifest(out, 'libri_%s_manifest.csv' % name, a.min_duration if is_train else None,
                            a.max_duration if is_train else None)
        print(name, n, 'utterances')


if __name__ == '__main__':
    main()
