"""Accuracy / loss curve PNGs from ``train.log`` + ``test.log``.

Parity target: ref:plot_curves.py:7-37 -- reads both logs with ``Logger.read``, writes
``test_accuracy.png`` (train blue / test red, legend lower right, title ``TEST accuracy``) and
``loss.png`` (legend upper right, title ``loss``) into ``save_path``.

Differences: rows may have 3 columns (reference format) or 4 (``epoch loss top1 top5``, the
format the reference README shows but its plotter cannot unpack -- SURVEY.md D7).  matplotlib is
used when importable; otherwise a small built-in rasteriser (numpy + zlib PNG encoder, 5x7
bitmap font) draws the same two figures, so the rank-0 plot step never needs an extra install.
"""
from __future__ import annotations

import os
import struct
import zlib
from typing import List, Sequence, Tuple

import numpy as np

from .meters import Logger

# ----------------------------------------------------------------------------- tiny rasteriser
_FONT = {
    '0': "01110 10001 10011 10101 11001 10001 01110", '1': "00100 01100 00100 00100 00100 00100 01110",
    '2': "01110 10001 00001 00010 00100 01000 11111", '3': "11110 00001 00001 01110 00001 00001 11110",
    '4': "00010 00110 01010 10010 11111 00010 00010", '5': "11111 10000 11110 00001 00001 10001 01110",
    '6': "00110 01000 10000 11110 10001 10001 01110", '7': "11111 00001 00010 00100 01000 01000 01000",
    '8': "01110 10001 10001 01110 10001 10001 01110", '9': "01110 10001 10001 01111 00001 00010 01100",
    '.': "00000 00000 00000 00000 00000 01100 01100", '-': "00000 00000 00000 11111 00000 00000 00000",
    ' ': "00000 00000 00000 00000 00000 00000 00000", '%': "11001 11010 00010 00100 01000 01011 10011",
    'A': "01110 10001 10001 11111 10001 10001 10001", 'B': "11110 10001 10001 11110 10001 10001 11110",
    'C': "01110 10001 10000 10000 10000 10001 01110", 'D': "11110 10001 10001 10001 10001 10001 11110",
    'E': "11111 10000 10000 11110 10000 10000 11111", 'F': "11111 10000 10000 11110 10000 10000 10000",
    'G': "01110 10001 10000 10111 10001 10001 01111", 'H': "10001 10001 10001 11111 10001 10001 10001",
    'I': "01110 00100 00100 00100 00100 00100 01110", 'J': "00111 00010 00010 00010 00010 10010 01100",
    'K': "10001 10010 10100 11000 10100 10010 10001", 'L': "10000 10000 10000 10000 10000 10000 11111",
    'M': "10001 11011 10101 10101 10001 10001 10001", 'N': "10001 11001 10101 10011 10001 10001 10001",
    'O': "01110 10001 10001 10001 10001 10001 01110", 'P': "11110 10001 10001 11110 10000 10000 10000",
    'Q': "01110 10001 10001 10001 10101 10010 01101", 'R': "11110 10001 10001 11110 10100 10010 10001",
    'S': "01111 10000 10000 01110 00001 00001 11110", 'T': "11111 00100 00100 00100 00100 00100 00100",
    'U': "10001 10001 10001 10001 10001 10001 01110", 'V': "10001 10001 10001 10001 10001 01010 00100",
    'W': "10001 10001 10001 10101 10101 11011 10001", 'X': "10001 10001 01010 00100 01010 10001 10001",
    'Y': "10001 10001 01010 00100 00100 00100 00100", 'Z': "11111 00001 00010 00100 01000 10000 11111",
}


class _Canvas:
    def __init__(self, w: int = 640, h: int = 480) -> None:
        self.w, self.h = w, h
        self.px = np.full((h, w, 3), 255, dtype=np.uint8)

    def dot(self, x: int, y: int, c) -> None:
        if 0 <= x < self.w and 0 <= y < self.h:
            self.px[y, x] = c

    def line(self, x0, y0, x1, y1, c, thick: int = 1) -> None:
        x0, y0, x1, y1 = int(round(x0)), int(round(y0)), int(round(x1)), int(round(y1))
        n = max(abs(x1 - x0), abs(y1 - y0), 1)
        for i in range(n + 1):
            x = x0 + (x1 - x0) * i // n
            y = y0 + (y1 - y0) * i // n
            for dx in range(thick):
                for dy in range(thick):
                    self.dot(x + dx, y + dy, c)

    def text(self, x: int, y: int, s: str, c=(0, 0, 0), scale: int = 2) -> None:
        for ch in s.upper():
            rows = _FONT.get(ch, _FONT[' ']).split()
            for r, bits in enumerate(rows):
                for q, b in enumerate(bits):
                    if b == '1':
                        for dx in range(scale):
                            for dy in range(scale):
                                self.dot(x + q * scale + dx, y + r * scale + dy, c)
            x += 6 * scale

    def save(self, path: str) -> None:
        raw = b''.join(b'\x00' + self.px[r].tobytes() for r in range(self.h))

        def chunk(tag: bytes, data: bytes) -> bytes:
            return (struct.pack('>I', len(data)) + tag + data +
                    struct.pack('>I', zlib.crc32(tag + data) & 0xffffffff))

        with open(path, 'wb') as f:
            f.write(b'\x89PNG\r\n\x1a\n')
            f.write(chunk(b'IHDR', struct.pack('>IIBBBBB', self.w, self.h, 8, 2, 0, 0, 0)))
            f.write(chunk(b'IDAT', zlib.compress(raw, 6)))
            f.write(chunk(b'IEND', b''))


def _nice_range(vals: Sequence[float]) -> Tuple[float, float]:
    lo, hi = float(min(vals)), float(max(vals))
    if hi - lo < 1e-12:
        lo, hi = lo - 0.5, hi + 0.5
    pad = 0.05 * (hi - lo)
    return lo - pad, hi + pad


def _raster_plot(path: str, x, series, ylabel: str, title: str, legend_loc: str) -> None:
    cv = _Canvas()
    L, R, T, B = 80, cv.w - 30, 50, cv.h - 60
    xlo, xhi = _nice_range(x)
    ylo, yhi = _nice_range([v for _, ys, _ in series for v in ys])
    cv.line(L, B, R, B, (0, 0, 0)); cv.line(L, T, L, B, (0, 0, 0))
    cv.line(R, T, R, B, (0, 0, 0)); cv.line(L, T, R, T, (0, 0, 0))
    for i in range(6):
        gx = L + (R - L) * i / 5
        gy = B - (B - T) * i / 5
        cv.line(gx, B, gx, B + 5, (0, 0, 0)); cv.line(L - 5, gy, L, gy, (0, 0, 0))
        cv.text(int(gx) - 12, B + 10, ('%.4g' % (xlo + (xhi - xlo) * i / 5))[:6], scale=1)
        cv.text(8, int(gy) - 4, ('%.4g' % (ylo + (yhi - ylo) * i / 5))[:8], scale=1)
    sx = lambda v: L + (v - xlo) / (xhi - xlo) * (R - L)
    sy = lambda v: B - (v - ylo) / (yhi - ylo) * (B - T)
    for _, ys, col in series:
        pts = [(sx(a), sy(b)) for a, b in zip(x, ys)]
        if len(pts) == 1:
            cv.line(pts[0][0] - 2, pts[0][1], pts[0][0] + 2, pts[0][1], col, 3)
        for (a0, b0), (a1, b1) in zip(pts[:-1], pts[1:]):
            cv.line(a0, b0, a1, b1, col, 2)
    cv.text((cv.w - 12 * len(title)) // 2, 18, title)
    cv.text((cv.w - 12 * 5) // 2, cv.h - 28, 'Epoch')
    cv.text(8, 28, ylabel, scale=1)
    lx = R - 120
    ly = (B - 20 - 18 * len(series)) if legend_loc.startswith('lower') else (T + 10)
    for i, (name, _, col) in enumerate(series):
        cv.line(lx, ly + 18 * i + 7, lx + 30, ly + 18 * i + 7, col, 2)
        cv.text(lx + 38, ly + 18 * i, name)
    cv.save(path)


def _plot(path: str, x, series, ylabel: str, title: str, legend_loc: str) -> str:
    try:
        import matplotlib
        matplotlib.use('Agg')
        import matplotlib.pyplot as plt
    except Exception:
        _raster_plot(path, x, series, ylabel, title, legend_loc)
        return 'builtin'
    for (name, ys, col), fmt in zip(series, ('-b', '-r')):
        plt.plot(x, ys, fmt, label=name)
    plt.xlabel('Epoch'); plt.ylabel(ylabel); plt.legend(loc=legend_loc); plt.title(title)
    plt.savefig(path); plt.close()
    return 'matplotlib'


def read_curves(save_path: str):
    """Returns ``(epochs, train_loss, train_acc, test_loss, test_acc)`` from the two logs.
    Accepts 3- or 4-column rows; the accuracy plotted is column 2 (top-1)."""
    tr = Logger(os.path.join(save_path, 'train.log')).read()
    te = Logger(os.path.join(save_path, 'test.log')).read()
    n = min(len(tr), len(te))
    tr, te = tr[:n], te[:n]
    epoch = [r[0] for r in tr]
    return (epoch, [r[1] for r in tr], [r[2] for r in tr], [r[1] for r in te], [r[2] for r in te])


def draw_plot(save_path: str) -> str:
    """Writes ``test_accuracy.png`` and ``loss.png`` (ref:plot_curves.py:19-37). Returns the
    renderer used (``'matplotlib'`` or ``'builtin'``)."""
    epoch, trl, tra, tel, tea = read_curves(save_path)
    blue, red = (31, 60, 220), (220, 40, 30)
    how = _plot(os.path.join(save_path, 'test_accuracy.png'), epoch,
                [('train', tra, blue), ('test', tea, red)], 'accuracy', 'TEST accuracy ', 'lower right')
    _plot(os.path.join(save_path, 'loss.png'), epoch,
          [('train', trl, blue), ('test', tel, red)], 'loss', 'loss', 'upper right')
    return how
