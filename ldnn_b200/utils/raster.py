"""Tiny dependency-free plotting backend: numpy canvas → PNG (zlib).

matplotlib is not part of this image, and the framework must still emit the
per-variant graphs; this module draws line charts and box plots well enough to read
(axes, ticks, legend, 5×7 bitmap text).  ``utils/viz.py`` prefers matplotlib when
it is importable and falls back to this.
"""
from __future__ import annotations

import struct
import zlib
from typing import List, Sequence

import numpy as np

_FONT = {
    "0": "01110 10001 10011 10101 11001 10001 01110", "1": "00100 01100 00100 00100 00100 00100 01110",
    "2": "01110 10001 00001 00010 00100 01000 11111", "3": "11110 00001 00001 01110 00001 00001 11110",
    "4": "00010 00110 01010 10010 11111 00010 00010", "5": "11111 10000 11110 00001 00001 10001 01110",
    "6": "00110 01000 10000 11110 10001 10001 01110", "7": "11111 00001 00010 00100 01000 01000 01000",
    "8": "01110 10001 10001 01110 10001 10001 01110", "9": "01110 10001 10001 01111 00001 00010 01100",
    ".": "00000 00000 00000 00000 00000 01100 01100", "-": "00000 00000 00000 11111 00000 00000 00000",
    "%": "11001 11010 00010 00100 01000 01011 10011", "/": "00001 00010 00010 00100 01000 01000 10000",
    ":": "00000 01100 01100 00000 01100 01100 00000", " ": "00000 00000 00000 00000 00000 00000 00000",
    "(": "00010 00100 01000 01000 01000 00100 00010", ")": "01000 00100 00010 00010 00010 00100 01000",
    "_": "00000 00000 00000 00000 00000 00000 11111", "+": "00000 00100 00100 11111 00100 00100 00000",
    "A": "01110 10001 10001 11111 10001 10001 10001", "B": "11110 10001 10001 11110 10001 10001 11110",
    "C": "01110 10001 10000 10000 10000 10001 01110", "D": "11110 10001 10001 10001 10001 10001 11110",
    "E": "11111 10000 10000 11110 10000 10000 11111", "F": "11111 10000 10000 11110 10000 10000 10000",
    "G": "01110 10001 10000 10111 10001 10001 01111", "H": "10001 10001 10001 11111 10001 10001 10001",
    "I": "01110 00100 00100 00100 00100 00100 01110", "J": "00111 00010 00010 00010 00010 10010 01100",
    "K": "10001 10010 10100 11000 10100 10010 10001", "L": "10000 10000 10000 10000 10000 10000 11111",
    "M": "10001 11011 10101 10101 10001 10001 10001", "N": "10001 11001 10101 10011 10001 10001 10001",
    "O": "01110 10001 10001 10001 10001 10001 01110", "P": "11110 10001 10001 11110 10000 10000 10000",
    "Q": "01110 10001 10001 10001 10101 10010 01101", "R": "11110 10001 10001 11110 10100 10010 10001",
    "S": "01111 10000 10000 01110 00001 00001 11110", "T": "11111 00100 00100 00100 00100 00100 00100",
    "U": "10001 10001 10001 10001 10001 10001 01110", "V": "10001 10001 10001 10001 10001 01010 00100",
    "W": "10001 10001 10001 10101 10101 11011 10001", "X": "10001 10001 01010 00100 01010 10001 10001",
    "Y": "10001 10001 01010 00100 00100 00100 00100", "Z": "11111 00001 00010 00100 01000 10000 11111",
}
_GLYPH = {k: np.array([[c == "1" for c in row] for row in v.split()], dtype=bool) for k, v in _FONT.items()}
COLORS = [(31, 119, 180), (255, 127, 14), (44, 160, 44), (214, 39, 40), (148, 103, 189), (140, 86, 75)]


class Canvas:
    def __init__(self, w=1600, h=1000):
        self.w, self.h = w, h
        self.px = np.full((h, w, 3), 255, dtype=np.uint8)

    def hline(self, x0, x1, y, c=(0, 0, 0)):
        if 0 <= y < self.h:
            self.px[y, max(0, min(x0, x1)):min(self.w, max(x0, x1) + 1)] = c

    def vline(self, x, y0, y1, c=(0, 0, 0)):
        if 0 <= x < self.w:
            self.px[max(0, min(y0, y1)):min(self.h, max(y0, y1) + 1), x] = c

    def line(self, x0, y0, x1, y1, c=(0, 0, 0), t=2):
        n = int(max(abs(x1 - x0), abs(y1 - y0))) + 1
        xs = np.linspace(x0, x1, n).round().astype(int)
        ys = np.linspace(y0, y1, n).round().astype(int)
        for dx in range(t):
            for dy in range(t):
                ok = (xs + dx >= 0) & (xs + dx < self.w) & (ys + dy >= 0) & (ys + dy < self.h)
                self.px[ys[ok] + dy, xs[ok] + dx] = c

    def rect(self, x0, y0, x1, y1, c=(0, 0, 0), fill=None):
        if fill is not None:
            self.px[max(0, min(y0, y1)):max(y0, y1) + 1, max(0, min(x0, x1)):max(x0, x1) + 1] = fill
        self.hline(x0, x1, y0, c); self.hline(x0, x1, y1, c); self.vline(x0, y0, y1, c); self.vline(x1, y0, y1, c)

    def dot(self, x, y, c, r=4):
        yy, xx = np.ogrid[-r:r + 1, -r:r + 1]
        m = xx * xx + yy * yy <= r * r
        y0, x0 = int(y) - r, int(x) - r
        if y0 < 0 or x0 < 0 or y0 + 2 * r + 1 > self.h or x0 + 2 * r + 1 > self.w:
            return
        self.px[y0:y0 + 2 * r + 1, x0:x0 + 2 * r + 1][m] = c

    def text(self, x, y, s, c=(0, 0, 0), scale=2):
        for ch in str(s).upper():
            g = _GLYPH.get(ch, _GLYPH[" "])
            gg = np.kron(g, np.ones((scale, scale), dtype=bool))
            hh, ww = gg.shape
            if 0 <= y and y + hh <= self.h and 0 <= x and x + ww <= self.w:
                self.px[y:y + hh, x:x + ww][gg] = c
            x += ww + scale

    def save(self, path):
        raw = b"".join(b"\x00" + self.px[r].tobytes() for r in range(self.h))

        def chunk(tag, data):
            return struct.pack(">I", len(data)) + tag + data + struct.pack(">I", zlib.crc32(tag + data) & 0xFFFFFFFF)

        with open(path, "wb") as f:
            f.write(b"\x89PNG\r\n\x1a\n" + chunk(b"IHDR", struct.pack(">IIBBBBB", self.w, self.h, 8, 2, 0, 0, 0))
                    + chunk(b"IDAT", zlib.compress(raw, 6)) + chunk(b"IEND", b""))


class Axes:
    def __init__(self, cv: Canvas, x0, y0, x1, y1, title="", xlabel="", ylabel=""):
        self.cv, self.box = cv, (x0, y0, x1, y1)
        cv.rect(x0, y0, x1, y1)
        cv.text(x0 + (x1 - x0) // 2 - 7 * len(title), y0 - 28, title, scale=3)
        cv.text(x0 + (x1 - x0) // 2 - 6 * len(xlabel), y1 + 44, xlabel)
        cv.text(max(x0 - 110, 2), y0 - 24, ylabel)
        self.xr, self.yr = (0.0, 1.0), (0.0, 1.0)

    def set_range(self, xr, yr):
        def pad(lo, hi):
            if not np.isfinite(lo) or not np.isfinite(hi):
                lo, hi = 0.0, 1.0
            if hi - lo < 1e-12:
                lo, hi = lo - 0.5, hi + 0.5
            m = 0.05 * (hi - lo)
            return lo - m, hi + m
        self.xr, self.yr = pad(*xr), pad(*yr)
        x0, y0, x1, y1 = self.box
        for i in range(6):
            v = self.yr[0] + (self.yr[1] - self.yr[0]) * i / 5
            py = self.py(v)
            self.cv.hline(x0, x1, py, (225, 225, 225))
            self.cv.text(max(x0 - 100, 2), py - 7, f"{v:.3g}")
        self.cv.rect(x0, y0, x1, y1)

    def px(self, v):
        x0, _, x1, _ = self.box
        return int(x0 + (v - self.xr[0]) / (self.xr[1] - self.xr[0]) * (x1 - x0))

    def py(self, v):
        _, y0, _, y1 = self.box
        return int(y1 - (v - self.yr[0]) / (self.yr[1] - self.yr[0]) * (y1 - y0))

    def xticks(self, positions, labels):
        _, _, _, y1 = self.box
        step = max(1, len(positions) // 25)
        for i, (p, l) in enumerate(zip(positions, labels)):
            if i % step:
                continue
            X = self.px(p)
            self.cv.vline(X, y1, y1 + 6)
            self.cv.text(X - 6 * len(str(l)), y1 + 12, l)

    def plot(self, xs, ys, color, label=None, slot=0):
        pts = [(self.px(x), self.py(y)) for x, y in zip(xs, ys) if np.isfinite(y)]
        for (a, b), (c, d) in zip(pts[:-1], pts[1:]):
            self.cv.line(a, b, c, d, color)
        for a, b in pts:
            self.cv.dot(a, b, color)
        if label:
            x0, y0, _, _ = self.box
            self.cv.line(x0 + 20, y0 + 20 + 24 * slot, x0 + 60, y0 + 20 + 24 * slot, color, 3)
            self.cv.text(x0 + 70, y0 + 13 + 24 * slot, label)

    def boxplot(self, groups: Sequence[Sequence[float]], labels: List[str]):
        flat = [v for g in groups for v in g if np.isfinite(v)]
        lo, hi = (min(flat), max(flat)) if flat else (0.0, 1.0)
        n = max(len(groups), 1)
        self.set_range((0.5, n + 0.5), (lo, hi))
        half = max(3, int(0.3 * (self.px(2) - self.px(1)))) if n > 1 else 60
        for i, g in enumerate(groups, start=1):
            a = np.asarray([v for v in g if np.isfinite(v)], dtype=np.float64)
            if a.size == 0:
                continue
            q1, med, q3 = np.percentile(a, [25, 50, 75])
            iqr = q3 - q1
            inl = a[(a >= q1 - 1.5 * iqr) & (a <= q3 + 1.5 * iqr)]
            wl, wh = (inl.min(), inl.max()) if inl.size else (q1, q3)
            X = self.px(i)
            self.cv.rect(X - half, self.py(q3), X + half, self.py(q1), (0, 0, 0), fill=(222, 235, 247))
            self.cv.line(X - half, self.py(med), X + half, self.py(med), COLORS[1], 3)
            self.cv.vline(X, self.py(wh), self.py(q3)); self.cv.vline(X, self.py(q1), self.py(wl))
            self.cv.hline(X - half // 2, X + half // 2, self.py(wh)); self.cv.hline(X - half // 2, X + half // 2, self.py(wl))
            for o in a[(a < wl) | (a > wh)][:200]:
                self.cv.dot(X, self.py(o), (90, 90, 90), 2)
        self.xticks(list(range(1, n + 1)), labels)
