#!/usr/bin/env python
"""Read one record back from the packed ImageNet file (the reference's scripts/test_read_hdf5.py opens a
matplotlib window; here the image is written to a PPM file instead)."""
import argparse

import numpy as np

if __name__ == '__main__':
    ap = argparse.ArgumentParser()
    ap.add_argument('--file', required=True)
    ap.add_argument('--index', type=int, default=0)
    ap.add_argument('--out', default='record.ppm')
    a = ap.parse_args()
    import h5py
    with h5py.File(a.file, 'r') as f:
        img = np.asarray(f['train_img'][a.index], dtype=np.uint8)
        label = int(f['train_labels'][a.index])
    with open(a.out, 'wb') as o:
        o.write(b'P6 %d %d 255\n' % (img.shape[1], img.shape[0]))
        o.write(img.tobytes())
    print('label', label, 'shape', img.shape, '->', a.out)
