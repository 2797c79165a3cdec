#!/usr/bin/env python
"""Pack an ImageNet folder tree into one HDF5 file (the reference's scripts/create_hdf5.py:1-114): datasets
``train_img`` / ``val_img`` uint8 [N,S,S,3] and ``train_labels`` / ``val_labels`` int16, shuffled.
Needs h5py and Pillow (the reference used cv2).

    python scripts/create_hdf5.py --root /data/ILSVRC2012 --out /tmp/imagenet_hdf5/imagenet-shuffled.hdf5
"""
import argparse
import os
import random

import numpy as np


def list_split(root, split, class_to_idx):
    items = []
    base = os.path.join(root, split)
    for cls in sorted(os.listdir(base)):
        d = os.path.join(base, cls)
        if os.path.isdir(d):
            items += [(os.path.join(d, f), class_to_idx[cls]) for f in sorted(os.listdir(d))]
    return items


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('--root', required=True)
    ap.add_argument('--out', required=True)
    ap.add_argument('--size', type=int, default=224)
    ap.add_argument('--seed', type=int, default=0)
    a = ap.parse_args()
    import h5py
    from PIL import Image
    classes = sorted(os.listdir(os.path.join(a.root, 'train')))
    class_to_idx = {c: i for i, c in enumerate(classes)}
    rng = random.Random(a.seed)
    # the class-directory -> label table next to the HDF5 file, one "wnid index" pair per line
    # (scripts/imagenet_label_mapping.csv of the reference is the output of create_hdf5.py:53-57)
    with open(os.path.join(os.path.dirname(os.path.abspath(a.out)), 'imagenet_label_mapping.csv'), 'w') as m:
        for c in classes:
            m.write('%s %d\n' % (c, class_to_idx[c]))
    with h5py.File(a.out, 'w') as f:
        for split in ('train', 'val'):
            items = list_split(a.root, split, class_to_idx)
            rng.shuffle(items)
            imgs = f.create_dataset(split + '_img', (len(items), a.size, a.size, 3), dtype=np.uint8)
            labels = f.create_dataset(split + '_labels', (len(items),), dtype=np.int16)
            for i, (path, y) in enumerate(items):
                im = Image.open(path).convert('RGB').resize((a.size, a.size), Image.BICUBIC)
                imgs[i] = np.asarray(im, dtype=np.uint8)
                labels[i] = y
                if i % 1000 == 0:
                    print('%s: %d/%d' % (split, i, len(items)))


if __name__ == '__main__':
    main()
