"""CPU: pins oracle/cv_spec.py (the executable specification csrc/mask_rect.cu and csrc/raster.cu implement)
against the installed OpenCV on randomised inputs.  OpenCV is an unpinned binary dependency of the reference
(requirements.txt:13); its version is recorded in the golden files."""
import cv2
import numpy as np
import pytest
from scipy import ndimage as ndi

import cv_spec as S


def _rand_hull(rs, W, H):
    """8 projected-corner-like points, clamped like spaces.py:240-243, hulled and truncated like scorefunction.py:101-102"""
    c = rs.uniform([-0.3 * W, -0.3 * H], [1.3 * W, 1.3 * H])
    pts = (c + rs.normal(0, rs.uniform(5, 0.6 * W), (8, 2))).astype(np.float32)
    pts[:, 0] = np.clip(pts[:, 0], int(-W / 2 + 1), 2 * W - 1)
    pts[:, 1] = np.clip(pts[:, 1], int(-H / 2 + 1), 2 * H - 1)
    return pts


def test_clip_line_matches_cv2():
    rs = np.random.RandomState(0)
    W, H = 97, 61
    for _ in range(20000):
        p1 = tuple(int(v) for v in rs.randint(-150, 250, 2))
        p2 = tuple(int(v) for v in rs.randint(-150, 250, 2))
        ok, q1, q2 = cv2.clipLine((0, 0, W, H), p1, p2)
        ok2, r1, r2 = S.clip_line(W, H, p1, p2)
        assert bool(ok) == ok2
        if ok:
            assert tuple(q1) == tuple(r1) and tuple(q2) == tuple(r2)


def test_fillpoly_row_intervals_match_cv2():
    rs = np.random.RandomState(1)
    W, H = 160, 120
    for t in range(250):
        pts = _rand_hull(rs, W, H)
        hull = cv2.convexHull(pts)
        poly = np.array([hull], dtype=np.int32)
        canvas = np.zeros((H, W), np.uint8)
        cv2.fillPoly(canvas, poly, 1)
        P = [tuple(int(v) for v in p) for p in poly[0, :, 0]]
        for y in range(0, H, 4 if t % 5 else 1):
            row = np.zeros(W, np.uint8)
            for xa, xb in S.fillpoly_row_intervals(W, H, P, y):
                row[max(xa, 0):min(xb, W - 1) + 1] = 1
            assert np.array_equal(row, canvas[y]), (t, y)


def test_strict_hull_matches_cv2_convexhull():
    rs = np.random.RandomState(2)
    for _ in range(3000):
        pts = _rand_hull(rs, 640, 480)
        hv = set(map(tuple, cv2.convexHull(pts)[:, 0].tolist()))
        assert hv == set((float(x), float(y)) for x, y in S.hull_strict_f32(pts))


def _blob(rs, H, W, kind):
    if kind == 0:
        return ndi.gaussian_filter(rs.randn(H, W), rs.uniform(2, 8)) > rs.uniform(-0.02, 0.08)
    if kind == 1:
        m = np.zeros((H, W), bool)
        for _ in range(rs.randint(1, 6)):
            x, y, s = rs.randint(0, W), rs.randint(0, H), rs.randint(1, 14)
            m[y:y + s, x:x + s] = True
        return m
    yy, xx = np.mgrid[0:H, 0:W]
    cx, cy = rs.uniform(30, W - 30), rs.uniform(30, H - 30)
    a, b = rs.uniform(8, 40, 2)
    th = rs.uniform(0, np.pi)
    u = (xx - cx) * np.cos(th) + (yy - cy) * np.sin(th)
    v = -(xx - cx) * np.sin(th) + (yy - cy) * np.cos(th)
    m = (u / a) ** 2 + (v / b) ** 2 <= 1
    if rs.rand() < 0.5:
        m &= ~((u / (0.5 * a)) ** 2 + (v / (0.5 * b)) ** 2 <= 1)      # ring: the hole must be filled
    return m


def test_contour_area_and_selection_match_cv2():
    rs = np.random.RandomState(3)
    H, W = 96, 128
    for t in range(120):
        m = _blob(rs, H, W, t % 3)
        contours, _ = cv2.findContours(m.astype(np.uint8), cv2.RETR_EXTERNAL, cv2.CHAIN_APPROX_SIMPLE)
        lab, area, n = S.external_component_areas(m)
        assert n == len(contours)
        if n == 0:
            continue
        assert sorted(cv2.contourArea(c) for c in contours) == sorted(area[1:n + 1].tolist())
        big = max(contours, key=cv2.contourArea)
        l = S.select_component(lab, area, n)
        x, y = big[0, 0]
        assert lab[y, x] == l


def test_mask_rect_spec_matches_cv2():
    rs = np.random.RandomState(4)
    H, W = 96, 128
    exact = close = 0
    for t in range(150):
        m = _blob(rs, H, W, t % 3)
        contours, _ = cv2.findContours(m.astype(np.uint8), cv2.RETR_EXTERNAL, cv2.CHAIN_APPROX_SIMPLE)
        r = S.mask_rect_spec(m)
        if not contours:
            assert r is None
            continue
        ref = cv2.boxPoints(cv2.minAreaRect(max(contours, key=cv2.contourArea)))
        dev = max(min(float(np.hypot(*(p - q))) for q in ref) for p in r)
        if sorted(map(tuple, r.tolist())) == sorted(map(tuple, ref.tolist())):
            exact += 1
        elif dev < 1e-3:
            close += 1
        else:
            pytest.fail("rectangle differs by %.3f px on blob %d" % (dev, t))
    assert exact > 100


def test_calipers_near_ties_match_cv2():
    """lattice images of rotated rectangles (near-parallel hull edges, near-tied candidate rectangles) at the C2 frame
    shape: the spec must reproduce cv2.minAreaRect + boxPoints, which pins the float / double split of
    cv_spec.rotating_calipers (a pure fp32 walk fails about a third of these)"""
    H, W = 530, 730
    rs = np.random.RandomState(5)
    yy, xx = np.mgrid[0:H, 0:W]
    exact = 0
    for t in range(40):
        cx, cy, th = rs.uniform(0.3, 0.7) * W, rs.uniform(0.3, 0.7) * H, rs.uniform(0, np.pi)
        u = (xx - cx) * np.cos(th) + (yy - cy) * np.sin(th)
        v = -(xx - cx) * np.sin(th) + (yy - cy) * np.cos(th)
        m = ((np.abs(u) < rs.uniform(3, 0.3 * W)) & (np.abs(v) < rs.uniform(1, 0.2 * H))).astype(np.uint8)
        contours, _ = cv2.findContours(m, cv2.RETR_EXTERNAL, cv2.CHAIN_APPROX_SIMPLE)
        ref = cv2.boxPoints(cv2.minAreaRect(max(contours, key=cv2.contourArea)))
        r = S.mask_rect_spec(m)
        dev = max(min(float(np.hypot(*(p - q))) for q in ref) for p in r)
        assert dev < 1e-3, (t, dev)
        exact += sorted(map(tuple, r.tolist())) == sorted(map(tuple, ref.tolist()))
    assert exact >= 36
