"""TEST INFRASTRUCTURE ONLY -- executable specification of the OpenCV behaviour the CUDA kernels restate.

OpenCV is an unpinned binary dependency of the reference (requirements.txt:13, used at
ProposalNetwork/scoring/scorefunction.py:62-68 and :101-103); its source is not under /root/reference.
These pure-Python restatements of its published algorithms are what csrc/mask_rect.cu and csrc/raster.cu
implement; tests/test_cv_spec.py checks them against the installed cv2 on randomised inputs, which is what
pins them.  Nothing in the product imports this file.
"""
import math

import numpy as np

f32 = np.float32
XY = 16
ONE = 1 << XY


# ------------------------------------------------------------------------------------------------------------
# cv2.fillPoly (shift=0, LINE_8, one contour) as a union of closed-form per-row intervals
# ------------------------------------------------------------------------------------------------------------
def _tdiv(a, b):
    q = abs(a) // abs(b)
    return q if (a >= 0) == (b > 0) else -q


def clip_line(W, H, p1, p2):
    """== cv2.clipLine((0,0,W,H), p1, p2): Cohen-Sutherland, y first, truncation toward zero"""
    (x1, y1), (x2, y2) = p1, p2
    r, b = W - 1, H - 1
    c1 = (x1 < 0) + (x1 > r) * 2 + (y1 < 0) * 4 + (y1 > b) * 8
    c2 = (x2 < 0) + (x2 > r) * 2 + (y2 < 0) * 4 + (y2 > b) * 8
    if (c1 & c2) == 0 and (c1 | c2) != 0:
        if c1 & 12:
            a = 0 if c1 < 8 else b
            x1 += int(float(a - y1) * (x2 - x1) / (y2 - y1)); y1 = a
            c1 = (x1 < 0) + (x1 > r) * 2
        if c2 & 12:
            a = 0 if c2 < 8 else b
            x2 += int(float(a - y2) * (x2 - x1) / (y2 - y1)); y2 = a
            c2 = (x2 < 0) + (x2 > r) * 2
        if (c1 & c2) == 0 and (c1 | c2) != 0:
            if c1:
                a = 0 if c1 == 1 else r
                y1 += int(float(a - x1) * (y2 - y1) / (x2 - x1)); x1 = a; c1 = 0
            if c2:
                a = 0 if c2 == 1 else r
                y2 += int(float(a - x2) * (y2 - y1) / (x2 - x1)); x2 = a; c2 = 0
    return (c1 | c2) == 0, (x1, y1), (x2, y2)


def line_row_extent(W, H, p0, p1, y):
    """x extent [xa, xb] of the pixels cv2.line(p0, p1, LINE_8) sets in row y, or None.  Closed form of the
    Bresenham recurrence err0 = major - 2 minor; minor steps taken before iteration k:
    j_k = floor((2 minor k + major - 1) / (2 major))."""
    inside = lambda p: 0 <= p[0] < W and 0 <= p[1] < H
    if not (inside(p0) and inside(p1)):
        ok, p0, p1 = clip_line(W, H, p0, p1)
        if not ok:
            return None
    (x0, y0), (x1, y1) = p0, p1
    if x1 < x0:
        x0, y0, x1, y1 = x1, y1, x0, y0
    dx, dy = x1 - x0, y1 - y0
    sy = 1 if dy >= 0 else -1
    dy = abs(dy)
    J = (y - y0) * sy
    if dy > dx:  # steep: one pixel per row
        major, minor = dy, dx
        if J < 0 or J > major:
            return None
        j = (2 * minor * J + major - 1) // (2 * major)
        return (x0 + j, x0 + j)
    major, minor = dx, dy
    if J < 0 or J > minor:
        return None
    if minor == 0:
        return (x0, x0 + major)
    num = 2 * major * J - major + 1
    kmin = 0 if num <= 0 else (num + 2 * minor - 1) // (2 * minor)
    kmax = min(major, (2 * major * (J + 1) - major) // (2 * minor))
    if kmin > kmax:
        return None
    return (x0 + kmin, x0 + kmax)


def scan_row(W, H, poly, y):
    """interior spans of fillPoly in row y (16.16 fixed point, x from the clipped edge copy, no +1/2 offset)"""
    xs = []
    for i in range(len(poly)):
        (x0, y0), (x1, y1) = poly[i - 1], poly[i]
        if y0 == y1 or not (min(y0, y1) <= y < max(y0, y1)):
            continue
        cx0, cy0, cx1, cy1 = x0, y0, x1, y1
        if not (0 <= x0 < W and 0 <= x1 < W and 0 <= y0 < H and 0 <= y1 < H):
            _, t0, t1 = clip_line(W, H, (x0, y0), (x1, y1))
            cx0, cx1 = t0[0], t1[0]
            if t0[1] != t1[1]:
                cy0, cy1 = t0[1], t1[1]
        X0, X1 = cx0 << XY, cx1 << XY
        dx = _tdiv(X1 - X0, cy1 - cy0)
        xs.append(X0 + (y0 - cy0) * dx + (y - y0) * dx if y0 < y1 else X1 + (y1 - cy1) * dx + (y - y1) * dx)
    xs.sort()
    out = []
    for k in range(0, len(xs) - 1, 2):
        xa, xb = (xs[k] + ONE - 1) >> XY, xs[k + 1] >> XY
        if xa < W and xb >= 0 and max(xa, 0) <= min(xb, W - 1):
            out.append((max(xa, 0), min(xb, W - 1)))
    return out


def fillpoly_row_intervals(W, H, poly, y):
    iv = scan_row(W, H, poly, y)
    for i in range(len(poly)):
        e = line_row_extent(W, H, poly[i - 1], poly[i], y)
        if e:
            iv.append(e)
    return iv


def hull_strict_f32(pts):
    """strict monotone chain on fp32 points (cross product of fp32 differences evaluated in double): the vertex
    set of cv2.convexHull, in the same rotational direction (image-clockwise)."""
    P = sorted(set((f32(x), f32(y)) for x, y in pts))
    if len(P) <= 2:
        return P

    def cross(o, a, b):
        ax, ay = float(f32(a[0] - o[0])), float(f32(a[1] - o[1]))
        bx, by = float(f32(b[0] - o[0])), float(f32(b[1] - o[1]))
        return ax * by - ay * bx
    lo = []
    for p in P:
        while len(lo) >= 2 and cross(lo[-2], lo[-1], p) <= 0:
            lo.pop()
        lo.append(p)
    up = []
    for p in reversed(P):
        while len(up) >= 2 and cross(up[-2], up[-1], p) <= 0:
            up.pop()
        up.append(p)
    return lo[:-1] + up[:-1]


# ------------------------------------------------------------------------------------------------------------
# findContours(RETR_EXTERNAL) -> max(contourArea): cell formula on the hole-filled 8-connected components
# ------------------------------------------------------------------------------------------------------------
def external_component_areas(mask):
    """Returns (labels [H,W] of the hole-filled 8-connected components, area per label).  contourArea of an
    external contour == sum over 2x2 pixel cells of {1: 4 corners set, 0.5: 3 corners set} of its filled
    component; filled components == 8-connected components of NOT(4-connected outside background)."""
    from scipy import ndimage as ndi
    H, W = mask.shape
    pad = np.zeros((H + 2, W + 2), bool)
    pad[1:-1, 1:-1] = mask
    bg, _ = ndi.label(~pad)
    M = ~(bg == bg[0, 0])
    lab, n = ndi.label(M, structure=np.ones((3, 3)))
    a = M[:-1, :-1].astype(int) + M[:-1, 1:] + M[1:, :-1] + M[1:, 1:]
    val = np.where(a == 4, 1.0, np.where(a == 3, 0.5, 0.0))
    l4 = np.maximum(np.maximum(lab[:-1, :-1], lab[:-1, 1:]), np.maximum(lab[1:, :-1], lab[1:, 1:]))
    area = np.bincount(l4.ravel(), weights=val.ravel(), minlength=n + 1)
    return lab[1:-1, 1:-1], area, n


def select_component(lab, area, n):
    """python max(contours, key=contourArea) on cv2's list order (last raster discovery first): the largest
    area wins, ties go to the component whose first pixel comes last in raster order."""
    first = {}
    ys, xs = np.nonzero(lab)
    for i in np.lexsort((xs, ys))[::-1]:
        first[lab[ys[i], xs[i]]] = (ys[i], xs[i])
    best = None
    for l in range(1, n + 1):
        if best is None or area[l] > area[best] or (area[l] == area[best] and first[l] > first[best]):
            best = l
    return best


# ------------------------------------------------------------------------------------------------------------
# cv2.minAreaRect + cv2.boxPoints
# ------------------------------------------------------------------------------------------------------------
def hull_int(pts):
    """strict hull of integer points in cv2.convexHull's output order: start at (max x, max y), then
    image-clockwise (down the right side, along the bottom, up the left side, along the top)."""
    P = sorted(set(map(tuple, pts)))
    if len(P) <= 2:
        return P[::-1]

    def cross(o, a, b):
        return (a[0] - o[0]) * (b[1] - o[1]) - (a[1] - o[1]) * (b[0] - o[0])
    lo = []
    for p in P:
        while len(lo) >= 2 and cross(lo[-2], lo[-1], p) <= 0:
            lo.pop()
        lo.append(p)
    up = []
    for p in reversed(P):
        while len(up) >= 2 and cross(up[-2], up[-1], p) <= 0:
            up.pop()
        up.append(p)
    return up[:-1] + lo[:-1]


def rotating_calipers(pts):
    """OpenCV rotatingCalipers(CALIPERS_MINAREARECT) as the installed cv2 4.13 evaluates it.  Established black-box
    (the OpenCV source is not in this image): of the 16 assignments of float / double to {edge vectors + inverse
    lengths, dot products + cosines, base vector, width / height / area} exactly one reproduces cv2.minAreaRect on
    722 hulls of seven mask families including lattice images of rotated rectangles, whose near-parallel hull edges
    make the choice of the advancing caliper sensitive to the last bits: inverse lengths and cosines in double, base
    vector and areas in float (tests/test_cv_spec.py::test_calipers_near_ties_match_cv2)."""
    n = len(pts)
    px = [f32(p[0]) for p in pts]
    py = [f32(p[1]) for p in pts]
    inv, vx, vy = [None] * n, [None] * n, [None] * n
    left = bottom = right = top = 0
    left_x = right_x = px[0]
    top_y = bottom_y = py[0]
    p0x, p0y = px[0], py[0]
    for i in range(n):
        if p0x < left_x: left_x, left = p0x, i
        if p0x > right_x: right_x, right = p0x, i
        if p0y > top_y: top_y, top = p0y, i
        if p0y < bottom_y: bottom_y, bottom = p0y, i
        j = i + 1 if i + 1 < n else 0
        dx, dy = float(px[j]) - float(p0x), float(py[j]) - float(p0y)
        vx[i], vy[i] = f32(dx), f32(dy)                     # differences of pixel coordinates: exact either way
        inv[i] = 1.0 / math.sqrt(dx * dx + dy * dy)         # kept in double (see the docstring)
        p0x, p0y = px[j], py[j]
    orientation = f32(0)
    ax, ay = float(vx[n - 1]), float(vy[n - 1])
    for i in range(n):
        bx, by = float(vx[i]), float(vy[i])
        conv = ax * by - ay * bx
        if conv != 0:
            orientation = f32(1.0) if conv > 0 else f32(-1.0)
            break
        ax, ay = bx, by
    base_a, base_b = orientation, f32(0)
    seq = [bottom, right, top, left]
    minarea = f32(np.finfo(np.float32).max)
    buf = None
    for _ in range(n):
        a, b = float(base_a), float(base_b)
        X = [float(vx[seq[i]]) for i in range(4)]
        Y = [float(vy[seq[i]]) for i in range(4)]
        dp = [a * X[0] + b * Y[0], -b * X[1] + a * Y[1], -a * X[2] - b * Y[2], b * X[3] - a * Y[3]]     # double
        maxcos, main = dp[0] * inv[seq[0]], 0
        for i in range(1, 4):
            c = dp[i] * inv[seq[i]]
            if c > maxcos:
                main, maxcos = i, c
        pi = seq[main]
        lead_x, lead_y = vx[pi] * f32(inv[pi]), vy[pi] * f32(inv[pi])                                  # float again
        if main == 0: base_a, base_b = lead_x, lead_y
        elif main == 1: base_a, base_b = lead_y, -lead_x
        elif main == 2: base_a, base_b = -lead_x, -lead_y
        else: base_a, base_b = -lead_y, lead_x
        seq[main] = 0 if seq[main] + 1 == n else seq[main] + 1
        dx, dy = px[seq[1]] - px[seq[3]], py[seq[1]] - py[seq[3]]
        width = dx * base_a + dy * base_b
        dx, dy = px[seq[2]] - px[seq[0]], py[seq[2]] - py[seq[0]]
        height = -dx * base_b + dy * base_a
        area = width * height
        if area <= minarea:
            minarea = area
            buf = (seq[3], base_a, width, base_b, height, seq[0])
    A1, B1, A2, B2 = buf[1], buf[3], -buf[3], buf[1]
    C1 = A1 * px[buf[0]] + py[buf[0]] * B1
    C2 = A2 * px[buf[5]] + py[buf[5]] * B2
    idet = f32(1.0) / (A1 * B2 - A2 * B1)
    return [(C1 * B2 - C2 * B1) * idet, (A1 * C2 - A2 * C1) * idet, A1 * buf[2], B1 * buf[2], A2 * buf[4], B2 * buf[4]]


def min_area_rect(pts_int):
    """== cv2.minAreaRect (4.13 angle convention: [-90, 0))"""
    h = hull_int(pts_int)
    n = len(h)
    if n > 2:
        out = rotating_calipers(h)
        cx = out[0] + (out[2] + out[4]) * f32(0.5)
        cy = out[1] + (out[3] + out[5]) * f32(0.5)
        w = f32(math.sqrt(float(out[2]) ** 2 + float(out[3]) ** 2))
        hh = f32(math.sqrt(float(out[4]) ** 2 + float(out[5]) ** 2))
        ad = math.atan2(float(out[3]), float(out[2])) * 180 / math.pi
        if ad >= 90:
            ang = f32(ad - 180)
        else:
            w, hh, ang = hh, w, f32(ad - 90)
        return (cx, cy), (w, hh), ang
    if n == 2:
        cx = (f32(h[0][0]) + f32(h[1][0])) * f32(0.5)
        cy = (f32(h[0][1]) + f32(h[1][1])) * f32(0.5)
        dx, dy = float(f32(h[1][0]) - f32(h[0][0])), float(f32(h[1][1]) - f32(h[0][1]))
        return (cx, cy), (f32(math.sqrt(dx * dx + dy * dy)), f32(0)), f32(float(f32(math.atan2(dy, dx))) * 180 / math.pi)
    return (f32(h[0][0]), f32(h[0][1])), (f32(0), f32(0)), f32(0)


def box_points(rect):
    """== cv2.boxPoints (RotatedRect::points)"""
    (cx, cy), (w, h), ang = rect
    _a = float(ang) * math.pi / 180.0
    b = f32(math.cos(_a)) * f32(0.5)
    a = f32(math.sin(_a)) * f32(0.5)
    p0 = (cx - a * h - b * w, cy + b * h - a * w)
    p1 = (cx + a * h - b * w, cy - b * h - a * w)
    return np.array([p0, p1, (f32(2) * cx - p0[0], f32(2) * cy - p0[1]), (f32(2) * cx - p1[0], f32(2) * cy - p1[1])],
                    dtype=np.float32)


def mask_rect_spec(mask):
    """the whole of scorefunction.py:62-68 without cv2: mask -> 4 rectangle points (or None)"""
    lab, area, n = external_component_areas(mask)
    if n == 0:
        return None
    l = select_component(lab, area, n)
    ys, xs = np.nonzero(lab == l)
    pts = set()
    for yy in np.unique(ys):
        r = xs[ys == yy]
        pts.add((int(r.min()), int(yy)))
        pts.add((int(r.max()), int(yy)))
    return box_points(min_area_rect(sorted(pts)))
