#!/usr/bin/env python
"""Distribution of dumped gradients (replacement of the reference's logs/plot_cdf.py; headless).

    GK_LOGGING_GRADIENTS=1 ./run.sh ...        # writes r<rank>_gradients_iter_<i>.npy
    python tools/plot_cdf.py --filename a.npy b.npy --legend iter0 iter100 --out grads.png
"""
import argparse
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from gaussiank_sgd_b200.utils.plotting import plot_gradient_distribution  # noqa: E402

if __name__ == '__main__':
    ap = argparse.ArgumentParser(description='gradient distribution plots')
    ap.add_argument('--filename', nargs='+', required=True)
    ap.add_argument('--legend', nargs='+', default=None)
    ap.add_argument('--out', default='plot.png')
    a = ap.parse_args()
    print(plot_gradient_distribution(a.filename, a.legend or a.filename, a.out))
