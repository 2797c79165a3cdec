#!/usr/bin/env python
"""Curves from training logs (replacement of the reference's logs/plot.py; headless, writes a file).

    python tools/plot.py --logs a.log b.log --labels topk gaussiank --kind accuracy --title resnet20 --out acc.png
"""
import argparse
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from gaussiank_sgd_b200.utils.plotting import plot_runs  # noqa: E402

if __name__ == '__main__':
    ap = argparse.ArgumentParser()
    ap.add_argument('--logs', nargs='+', required=True)
    ap.add_argument('--labels', nargs='+', default=None)
    ap.add_argument('--kind', default='accuracy', choices=['accuracy', 'top5', 'top1', 'loss', 'lr', 'speed', 'latency'])
    ap.add_argument('--title', default='')
    ap.add_argument('--out', default='plot.png')
    a = ap.parse_args()
    print(plot_runs(a.logs, a.labels or [os.path.basename(l) for l in a.logs], a.kind, a.title, a.out))
