"""Minimal dependency-free SVG drawing used by the headless plotters."""
from __future__ import annotations

import numpy

PALETTE = ["#1f77b4", "#ff7f0e", "#2ca02c", "#d62728", "#9467bd", "#8c564b", "#e377c2", "#7f7f7f"]


def _esc(s):
    return str(s).replace("&", "&amp;").replace("<", "&lt;").replace(">", "&gt;")


class Canvas(object):
    def __init__(self, width=640, height=420, title=""):
        self.w, self.h = width, height
        self.items = ['<rect width="%d" height="%d" fill="white"/>' % (width, height)]
        if title:
            self.text(width / 2, 18, title, size=14, anchor="middle")
        self.plot = (56, 30, width - 20, height - 40)   # left, top, right, bottom

    def text(self, x, y, s, size=11, anchor="start", color="#222"):
        self.items.append('<text x="%.1f" y="%.1f" font-size="%d" font-family="sans-serif" text-anchor="%s" '
                          'fill="%s">%s</text>' % (x, y, size, anchor, color, _esc(s)))

    def axes(self, xmin, xmax, ymin, ymax, xlabel="", ylabel=""):
        l, t, r, b = self.plot
        # degenerate or huge ranges (a curve of identical values, 1e30 "not measured yet" markers)
        if not (xmax - xmin > 1e-12 * max(abs(xmin), abs(xmax), 1.0)):
            xmax = xmin + max(1.0, abs(xmin) * 1e-6)
        if not (ymax - ymin > 1e-12 * max(abs(ymin), abs(ymax), 1.0)):
            ymax = ymin + max(1.0, abs(ymin) * 1e-6)
        self.range = (xmin, xmax, ymin, ymax)
        self.items.append('<rect x="%d" y="%d" width="%d" height="%d" fill="none" stroke="#444"/>'
                          % (l, t, r - l, b - t))
        for i in range(5):
            fx = i / 4.0
            x = l + fx * (r - l)
            y = b - fx * (b - t)
            self.text(x, b + 14, "%.4g" % (xmin + fx * (xmax - xmin)), anchor="middle")
            self.text(l - 4, y + 4, "%.4g" % (ymin + fx * (ymax - ymin)), anchor="end")
            self.items.append('<line x1="%.1f" y1="%d" x2="%.1f" y2="%d" stroke="#ddd"/>' % (x, t, x, b))
            self.items.append('<line x1="%d" y1="%.1f" x2="%d" y2="%.1f" stroke="#ddd"/>' % (l, y, r, y))
        if xlabel:
            self.text((l + r) / 2, self.h - 6, xlabel, anchor="middle")
        if ylabel:
            self.text(12, t - 8, ylabel)

    def _xy(self, x, y):
        l, t, r, b = self.plot
        xmin, xmax, ymin, ymax = self.range
        return l + (x - xmin) / (xmax - xmin) * (r - l), b - (y - ymin) / (ymax - ymin) * (b - t)

    def line(self, xs, ys, color=None, label=None, index=0):
        color = color or PALETTE[index % len(PALETTE)]
        pts = " ".join("%.1f,%.1f" % self._xy(x, y) for x, y in zip(xs, ys))
        self.items.append('<polyline points="%s" fill="none" stroke="%s" stroke-width="1.5"/>' % (pts, color))
        if label:
            self.text(self.plot[2] - 6, self.plot[1] + 14 + 14 * index, label, anchor="end", color=color)

    def bars(self, edges, counts, color="#1f77b4"):
        for i, c in enumerate(counts):
            x0, y0 = self._xy(edges[i], c)
            x1, y1 = self._xy(edges[i + 1], 0)
            self.items.append('<rect x="%.1f" y="%.1f" width="%.1f" height="%.1f" fill="%s" stroke="white"/>'
                              % (x0, y0, max(x1 - x0, 0.5), max(y1 - y0, 0), color))

    def matrix(self, m, labels=None, annotate=True):
        m = numpy.asarray(m, dtype=numpy.float64)
        l, t, r, b = self.plot
        rows, cols = m.shape
        cw, ch = (r - l) / cols, (b - t) / rows
        top = m.max() if m.size and m.max() > 0 else 1.0
        for i in range(rows):
            for j in range(cols):
                v = m[i, j] / top
                shade = int(255 - 200 * v)
                self.items.append('<rect x="%.1f" y="%.1f" width="%.1f" height="%.1f" fill="rgb(%d,%d,255)" '
                                  'stroke="#999"/>' % (l + j * cw, t + i * ch, cw, ch, shade, shade))
                if annotate and rows <= 20:
                    self.text(l + (j + 0.5) * cw, t + (i + 0.6) * ch, "%g" % m[i, j], size=9, anchor="middle")
        if labels:
            for j, lab in enumerate(labels[:cols]):
                self.text(l + (j + 0.5) * cw, b + 14, lab, size=9, anchor="middle")

    def image(self, arr, x, y, w, h):
        """Embed a grayscale/RGB array as a grid of rects (small images only) — keeps the SVG dependency-free."""
        a = numpy.asarray(arr, dtype=numpy.float64)
        if a.ndim == 3 and a.shape[2] == 1:
            a = a[:, :, 0]
        lo, hi = a.min(), a.max()
        a = (a - lo) / (hi - lo) * 255 if hi > lo else numpy.zeros_like(a)
        step = max(1, int(max(a.shape[:2]) / 48))
        a = a[::step, ::step]
        rows, cols = a.shape[:2]
        pw, ph = w / cols, h / rows
        for i in range(rows):
            for j in range(cols):
                px = a[i, j]
                rgb = (int(px),) * 3 if numpy.ndim(px) == 0 else tuple(int(v) for v in px[:3])
                self.items.append('<rect x="%.1f" y="%.1f" width="%.2f" height="%.2f" fill="rgb(%d,%d,%d)"/>'
                                  % ((x + j * pw, y + i * ph, pw + 0.3, ph + 0.3) + rgb))

    def render(self):
        return ('<svg xmlns="http://www.w3.org/2000/svg" width="%d" height="%d" viewBox="0 0 %d %d">\n%s\n</svg>\n'
                % (self.w, self.h, self.w, self.h, "\n".join(self.items)))

    def save(self, path):
        with open(path, "w") as f:
            f.write(self.render())
        return path
