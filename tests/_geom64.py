"""Independent float64 cross-check (test helper): Sutherland-Hodgman clip +
shoelace in pure Python.  Shares no code and no algorithm with the oracle
(edge/edge hits + Graham scan) — disagreement between the two is a bug in the
oracle restatement."""
import math


def corners(b):
    cx, cy, w, h, a = (float(v) for v in b[:5])
    c, s = math.cos(a), math.sin(a)
    out = []
    for sx, sy in ((-1, -1), (1, -1), (1, 1), (-1, 1)):   # counter-clockwise
        x, y = sx * w / 2, sy * h / 2
        out.append((cx + c * x - s * y, cy + s * x + c * y))
    return out


def _clip(poly, a, b):
    """keep the part of poly on the left of directed line a->b"""
    out = []
    n = len(poly)
    for i in range(n):
        p, q = poly[i], poly[(i + 1) % n]
        sp = (b[0] - a[0]) * (p[1] - a[1]) - (b[1] - a[1]) * (p[0] - a[0])
        sq = (b[0] - a[0]) * (q[1] - a[1]) - (b[1] - a[1]) * (q[0] - a[0])
        if sp >= 0:
            out.append(p)
        if (sp > 0 and sq < 0) or (sp < 0 and sq > 0):
            t = sp / (sp - sq)
            out.append((p[0] + t * (q[0] - p[0]), p[1] + t * (q[1] - p[1])))
    return out


def inter_area(b1, b2):
    poly = corners(b1)
    clipper = corners(b2)
    for i in range(4):
        if not poly:
            return 0.0
        poly = _clip(poly, clipper[i], clipper[(i + 1) % 4])
    if len(poly) < 3:
        return 0.0
    s = 0.0
    for i in range(len(poly)):
        x0, y0 = poly[i]
        x1, y1 = poly[(i + 1) % len(poly)]
        s += x0 * y1 - x1 * y0
    return abs(s) / 2


def iou(b1, b2, mode='iou'):
    a1, a2 = float(b1[2]) * float(b1[3]), float(b2[2]) * float(b2[3])
    it = inter_area(b1, b2)
    return it / (a1 + a2 - it) if mode == 'iou' else it / a1
