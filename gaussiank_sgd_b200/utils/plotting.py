"""Log parsing and headless plotting (the reference's logs/plot.py, logs/plot_cdf.py, logs/utils.py).

The reference plots through the interactive TkAgg backend (logs/plot.py:11); there is no display here, so
figures go to files: matplotlib's Agg backend when matplotlib is installed, otherwise a small built-in SVG
writer.  The parsed log lines are the ones the trainer emits (dl_trainer.py:822, dist_trainer.py:90):
``Epoch E, lr: L, val loss: X, val top-1 acc: A, top-5 acc: B`` and ``Time per iteration ... Speed: S images/s``.
"""
from __future__ import annotations

import datetime
import re
from typing import Dict, List, Sequence

import numpy as np

STANDARD_TITLES = {'resnet20': 'ResNet-20', 'vgg16': 'VGG-16', 'alexnet': 'AlexNet', 'resnet50': 'ResNet-50',
                   'lstm': 'LSTM-PTB', 'lstmptb': 'LSTM-PTB', 'lstman4': 'LSTM-AN4'}
_EVAL = re.compile(r'Epoch (\d+), lr: ([0-9.eE+-]+), val loss: ([0-9.eE+-]+|nan|inf), '
                   r'val top-1 acc: ([0-9.eE+-]+|nan|inf), top-5 acc: ([0-9.eE+-]+|nan|inf)')
_SPEED = re.compile(r'Time per iteration including communication: ([0-9.eE+-]+)[,.] Speed: ([0-9.eE+-]+) images/s')
_STAMP = re.compile(r'^(\d{4}-\d{2}-\d{2} \d{2}:\d{2}:\d{2})')


def get_real_title(title: str) -> str:
    return STANDARD_TITLES.get(title, title)


def seconds_between_datetimestring(a: str, b: str) -> int:
    fmt = '%Y-%m-%d %H:%M:%S'
    d = datetime.datetime.strptime(b, fmt) - datetime.datetime.strptime(a, fmt)
    return d.days * 86400 + d.seconds


def read_data_from_log(logfile: str) -> Dict[str, List[float]]:
    """Per-epoch validation metrics, learning rates, wall-clock offsets and iteration speeds of one run."""
    out = {'epoch': [], 'lr': [], 'val_loss': [], 'top1': [], 'top5': [], 'seconds': [], 'iter_time': [], 'speed': []}
    t0 = None
    with open(logfile) as f:
        for line in f:
            stamp = _STAMP.match(line)
            m = _EVAL.search(line)
            if m:
                out['epoch'].append(int(m.group(1)))
                for key, g in (('lr', 2), ('val_loss', 3), ('top1', 4), ('top5', 5)):
                    out[key].append(float(m.group(g)))
                if stamp:
                    t0 = t0 or stamp.group(1)
                    out['seconds'].append(seconds_between_datetimestring(t0, stamp.group(1)))
            m = _SPEED.search(line)
            if m:
                out['iter_time'].append(float(m.group(1)))
                out['speed'].append(float(m.group(2)))
    return out


def _svg_lines(series: Sequence, labels: Sequence[str], title: str, xlabel: str, ylabel: str, path: str,
               width=640, height=420):
    """Minimal dependency-free line plot."""
    colors = ['#1f77b4', '#d62728', '#2ca02c', '#ff7f0e', '#9467bd', '#8c564b', '#17becf', '#7f7f7f']
    pts = [(np.asarray(x, float), np.asarray(y, float)) for x, y in series if len(x)]
    if not pts:
        raise ValueError('nothing to plot')
    xmin, xmax = min(p[0].min() for p in pts), max(p[0].max() for p in pts)
    ymin, ymax = min(p[1].min() for p in pts), max(p[1].max() for p in pts)
    xmax, ymax = (xmax if xmax > xmin else xmin + 1), (ymax if ymax > ymin else ymin + 1)
    L, R, Tp, B = 70, 20, 40, 50
    sx = lambda v: L + (v - xmin) / (xmax - xmin) * (width - L - R)    # noqa: E731
    sy = lambda v: height - B - (v - ymin) / (ymax - ymin) * (height - Tp - B)   # noqa: E731
    el = ['<svg xmlns="http://www.w3.org/2000/svg" width="%d" height="%d" font-family="sans-serif" font-size="12">' % (width, height),
          '<rect width="100%" height="100%" fill="white"/>',
          '<text x="%d" y="20" text-anchor="middle" font-size="15">%s</text>' % (width // 2, title),
          '<line x1="%d" y1="%d" x2="%d" y2="%d" stroke="black"/>' % (L, height - B, width - R, height - B),
          '<line x1="%d" y1="%d" x2="%d" y2="%d" stroke="black"/>' % (L, Tp, L, height - B),
          '<text x="%d" y="%d" text-anchor="middle">%s</text>' % (width // 2, height - 10, xlabel),
          '<text x="15" y="%d" text-anchor="middle" transform="rotate(-90 15 %d)">%s</text>' % (height // 2, height // 2, ylabel)]
    for i in range(5):
        xv, yv = xmin + (xmax - xmin) * i / 4, ymin + (ymax - ymin) * i / 4
        el.append('<text x="%.1f" y="%d" text-anchor="middle">%.4g</text>' % (sx(xv), height - B + 16, xv))
        el.append('<text x="%d" y="%.1f" text-anchor="end">%.4g</text>' % (L - 6, sy(yv) + 4, yv))
    for i, ((x, y), lab) in enumerate(zip(pts, labels)):
        c = colors[i % len(colors)]
        el.append('<polyline fill="none" stroke="%s" stroke-width="1.5" points="%s"/>' % (
            c, ' '.join('%.1f,%.1f' % (sx(a), sy(b)) for a, b in zip(x, y))))
        el.append('<text x="%d" y="%d" fill="%s">%s</text>' % (width - R - 150, Tp + 14 * (i + 1), c, lab))
    el.append('</svg>')
    with open(path, 'w') as f:
        f.write('\n'.join(el))
    return path


def _draw(series, labels, title, xlabel, ylabel, path):
    try:
        import matplotlib
        matplotlib.use('Agg')          # headless: files only, never a window
        import matplotlib.pyplot as plt
    except ImportError:
        if not path.endswith('.svg'):
            path = path.rsplit('.', 1)[0] + '.svg'
        return _svg_lines(series, labels, title, xlabel, ylabel, path)
    fig, ax = plt.subplots(1, 1, figsize=(5, 3.5))
    for (x, y), lab in zip(series, labels):
        ax.plot(x, y, label=lab, linewidth=1.5)
    ax.set_xlabel(xlabel); ax.set_ylabel(ylabel); ax.set_title(title); ax.grid(linestyle=':'); ax.legend(fontsize=8)
    fig.tight_layout()
    fig.savefig(path)
    plt.close(fig)
    return path


def plot_runs(logfiles: Sequence[str], labels: Sequence[str], kind: str = 'accuracy', title: str = '', out: str = 'plot.png'):
    """kind in {accuracy (top-5, what the reference's parser picks up, logs/plot.py:51-59), top1, loss, lr, speed,
    latency (metric over wall-clock)}."""
    series = []
    for lf in logfiles:
        d = read_data_from_log(lf)
        if kind in ('accuracy', 'top5'):
            series.append((d['epoch'], d['top5']))
        elif kind == 'top1':
            series.append((d['epoch'], d['top1']))
        elif kind == 'loss':
            series.append((d['epoch'], d['val_loss']))
        elif kind == 'lr':
            series.append((d['epoch'], d['lr']))
        elif kind == 'speed':
            series.append((list(range(len(d['speed']))), d['speed']))
        elif kind == 'latency':
            series.append((d['seconds'], d['top5'][:len(d['seconds'])]))
        else:
            raise ValueError(kind)
    ylabel = {'accuracy': 'top-5 accuracy [%]', 'top5': 'top-5 accuracy [%]', 'top1': 'top-1 accuracy [%]',
              'loss': 'validation loss', 'lr': 'learning rate', 'speed': 'images/s', 'latency': 'top-5 accuracy [%]'}[kind]
    xlabel = 'seconds' if kind == 'latency' else ('log sample' if kind == 'speed' else 'epoch')
    return _draw(series, labels, get_real_title(title), xlabel, ylabel, out)


def plot_gradient_distribution(files: Sequence[str], labels: Sequence[str], out: str = 'plot.png', bins: int = 200):
    """Density of saved gradient vectors (.npy dumps of GK_LOGGING_GRADIENTS=1; logs/plot_cdf.py:18-33 draws a
    KDE with seaborn, a normalised histogram is drawn here)."""
    series = []
    for f in files:
        x = np.load(f, allow_pickle=True).astype(np.float64).ravel()
        hist, edges = np.histogram(x, bins=bins, density=True)
        series.append((0.5 * (edges[1:] + edges[:-1]), hist))
        print('Name: %s Min: %s Max: %s' % (f.split('/')[-1], x.min(), x.max()))
    return _draw(series, labels, 'gradient distribution', 'gradient value', 'density', out)
