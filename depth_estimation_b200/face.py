"""Host-side scalar geometry of the reference's `face.py` (`/root/reference/face.py:4-168`): same class name,
attributes, method names and arithmetic - including the indexing quirks of `get_depth` (`:91-116`), which are
reproduced, not fixed, so the metric tail stays bit-identical (golden vectors generated from the reference
file itself: tests/golden/face_tail.npz).  A handful of scalar flops per frame - deliberately not on the GPU;
the dense part of the tail (whole-map /2.54, iris sampling) is fused into K8 (`ops.postprocess`).
"""
from __future__ import annotations

import math

import numpy as np

CARD_WIDTH_MM = 85.6      # face.py:17
IRIS_WIDTH_MM = 11.7      # face.py:19


class FaceDet(object):
    def __init__(self, d_2_obj, points):
        self.w_card = CARD_WIDTH_MM
        self.w_iris = IRIS_WIDTH_MM
        self.w_pix = math.dist(points[:2], points[2:])               # face.py:22
        self.d_2_obj = self.in_to_mm(d_2_obj)                        # face.py:24
        self.f = self.f_length()                                     # face.py:26
        self.l_iris = {'center': None, 'radius': None}
        self.r_iris = {'center': None, 'radius': None}
        self.mesh = None
        self.head_pts = None
        self.head_w = 0
        self.head_measurements = []
        self.s2c_d = 0
        self.s2c_ds = []
        self.ri_depth = 0
        self.ri_depths = []
        self.abs_depth = 0
        self.abs_depths = []
        self.error = 0
        self.errors = []

    # ---- camera geometry (face.py:51-89)
    def f_length(self):
        return (self.d_2_obj * self.w_pix) / self.w_card

    def s2c_dist(self, w_object, w_pix):
        cm = ((self.f * w_object) / w_pix) / 10
        self.s2c_ds.append(cm)
        self.s2c_d = self.cm_to_ft(cm)

    def get_headw(self, p1, p2):
        self.head_pts = (p1, p2)
        pix = math.dist((p1[0], p1[1]), (p2[0], p2[1]))
        self.head_w = (pix * self.w_iris) / (self.l_iris['radius'] * 2)
        self.head_measurements.append(self.head_w)

    # ---- metric tail (face.py:91-116, 132-139)
    def sample_indices(self, shape):
        """The two (i, j) index pairs `get_depth` reads from a map of `shape` - the reference's arithmetic:
        centres are (x, y); element 0 is clamped to shape[0]-1; for element 1 the LEFT entry is assigned twice
        (the second time from the right iris) and the right entry is never clamped; the map is indexed [x, y]."""
        if self.mesh is not None:
            l = [int(v) for v in self.l_iris['center']]
            r = [int(v) for v in self.r_iris['center']]
            l0, r0 = min(shape[0] - 1, l[0]), min(shape[0] - 1, r[0])
            l1 = min(shape[1] - 1, r[1])
            return (l0, l1), (r0, r[1])
        return (self.head_pts[0][0], self.head_pts[0][1]), (self.head_pts[1][0], self.head_pts[1][1])

    def get_depth(self, img):
        a, b = self.sample_indices(img.shape)
        ri_depth = (img[a[0], a[1]] + img[b[0], b[1]]) / 2
        self.ri_depth = ri_depth
        self.ri_depths.append(ri_depth)

    def set_depth_samples(self, left, right):
        """Same bookkeeping as get_depth when the two samples come from K8's fused gather."""
        ri_depth = (left + right) / 2
        self.ri_depth = ri_depth
        self.ri_depths.append(ri_depth)

    def rel2abs_2(self, pred_depths, gt_depths):
        gt = [1 / x for x in gt_depths]
        A = np.vstack([gt, np.ones(len(gt))]).T
        self.m, self.b = np.linalg.lstsq(A, pred_depths, rcond=None)[0]
        self.abs_depth = self.cm_to_ft(self.ri_depth * self.m + self.b)

    def rel2abs(self):
        abs_depth = self.ri_depth / 2.54
        self.abs_depth = abs_depth
        self.abs_depths.append(abs_depth)

    # ---- diagnostics (face.py:141-153)
    def rmse(self):
        sq = [(a - b) ** 2 for a, b in zip(self.s2c_ds, self.abs_depths)]
        return math.sqrt(sum(sq) / len(sq))

    def mae(self):
        ab = [abs(a - b) for a, b in zip(self.s2c_ds, self.abs_depths)]
        return sum(ab) / len(ab)

    # ---- unit helpers (face.py:155-168)
    def mm2cm(self, dist):
        return dist / 10

    def cm_to_ft(self, dist):
        return round(dist / (2.54 * 12), 2)

    def in_to_mm(self, dist):
        return round(dist * 2.54 * 10, 2)

    def diameter(self, radius):
        return int(radius * 2)

    def dist_euclid(self, pt1: tuple, pt2: tuple):
        return math.sqrt((pt2[0] - pt1[0]) ** 2 + (pt2[1] - pt1[1]) ** 2)
