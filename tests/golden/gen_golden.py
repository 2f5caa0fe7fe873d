#!/usr/bin/env python3
"""Generate the golden fixtures under tests/golden/ with python cv2 (OpenCV 4.13.0).

This is the INDEPENDENT pin of the CPU oracle: every pixel primitive is the real
OpenCV call the reference makes (cv2.resize, cv2.FastFeatureDetector, cv2.GaussianBlur,
cv2.fastAtan2), glued together by a small pure-python reading of
src/ORBextractor.cc (cell loop :776-850, quadtree :481-765 with the creation-order
tie rule of SURVEY.md B.1, IC_Angle :77-104, computeOrbDescriptor :108-147).
It cannot travel to the GPU box (cv2 may be absent / a different version there), so
its outputs are committed as small .npz files and the tests compare the oracle
(and, through it, the CUDA path) with them.

    python tests/golden/gen_golden.py        # rewrites tests/golden/*.npz
"""
import math
import os
import sys

import cv2
import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
from my_orb_slam2_b200 import synth  # noqa: E402


def load_pattern():
    txt = open(os.path.join(ROOT, "include", "orb_pattern_31.inc")).read()
    txt = txt[txt.index("*/") + 2:]
    return np.array([int(v) for v in txt.replace("\n", "").split(",") if v.strip()], np.int32).reshape(512, 2)


def f32(x):
    return np.float32(x)


def cv_round(x):
    return int(np.rint(np.float32(x)))  # round half to even


class PyExtractor:
    def __init__(self, nfeatures, scale_factor, nlevels, ini_th, min_th):
        self.nfeatures, self.nlevels, self.ini_th, self.min_th = nfeatures, nlevels, ini_th, min_th
        sf = float(f32(scale_factor))
        self.scale = [f32(1.0)]
        for i in range(1, nlevels):
            self.scale.append(f32(float(self.scale[-1]) * sf))
        self.inv_scale = [f32(1.0) / s for s in self.scale]
        factor = f32(1.0 / sf)
        desired = f32(nfeatures) * (f32(1) - factor) / (f32(1) - f32(math.pow(float(factor), float(nlevels))))
        self.quota, tot = [], 0
        for _ in range(nlevels - 1):
            self.quota.append(cv_round(desired))
            tot += self.quota[-1]
            desired = f32(desired * factor)
        self.quota.append(max(nfeatures - tot, 0))
        self.pattern = load_pattern()
        umax = [0] * 16
        vmax = int(math.floor(15 * math.sqrt(2.0) / 2 + 1))
        vmin = int(math.ceil(15 * math.sqrt(2.0) / 2))
        for v in range(vmax + 1):
            umax[v] = int(np.rint(math.sqrt(225.0 - v * v)))
        v0 = 0
        for v in range(15, vmin - 1, -1):
            while umax[v0] == umax[v0 + 1]:
                v0 += 1
            umax[v] = v0
            v0 += 1
        self.umax = umax

    def pyramid(self, img):
        levels = []
        for lv in range(self.nlevels):
            w = cv_round(f32(img.shape[1]) * self.inv_scale[lv])
            h = cv_round(f32(img.shape[0]) * self.inv_scale[lv])
            if lv == 0:
                levels.append(img.copy())
            else:
                levels.append(cv2.resize(levels[-1], (w, h), interpolation=cv2.INTER_LINEAR))
        return levels

    def cell_candidates(self, im):
        h, w = im.shape
        min_b, max_bx, max_by = 16, w - 16, h - 16
        width, height = f32(max_bx - min_b), f32(max_by - min_b)
        n_cols, n_rows = int(width / f32(30)), int(height / f32(30))
        w_cell, h_cell = int(math.ceil(width / f32(n_cols))), int(math.ceil(height / f32(n_rows)))
        det_hi = cv2.FastFeatureDetector_create(self.ini_th, True, cv2.FAST_FEATURE_DETECTOR_TYPE_9_16)
        det_lo = cv2.FastFeatureDetector_create(self.min_th, True, cv2.FAST_FEATURE_DETECTOR_TYPE_9_16)
        out = []
        for i in range(n_rows):
            ini_y = min_b + i * h_cell
            max_y = ini_y + h_cell + 6
            if ini_y >= max_by - 3:
                continue
            max_y = min(max_y, max_by)
            for j in range(n_cols):
                ini_x = min_b + j * w_cell
                max_x = ini_x + w_cell + 6
                if ini_x >= max_bx - 6:
                    continue
                max_x = min(max_x, max_bx)
                roi = np.ascontiguousarray(im[ini_y:max_y, ini_x:max_x])
                kps = det_hi.detect(roi)
                if len(kps) == 0:
                    kps = det_lo.detect(roi)
                for k in kps:
                    out.append((int(k.pt[0]) + j * w_cell, int(k.pt[1]) + i * h_cell, int(k.response)))
        return out

    @staticmethod
    def quadtree(cands, min_x, max_x, min_y, max_y, n_target):
        """Pure-python reading of DistributeOctTree; nodes are dicts, list order as std::list."""
        if not cands:
            return []
        n_ini = int(np.round(f32(max_x - min_x) / f32(max_y - min_y)))
        hx = f32(max_x - min_x) / f32(n_ini)
        seq = [0]

        def mk(x0, x1, y0, y1, keys):
            seq[0] += 1
            return dict(x0=x0, x1=x1, y0=y0, y1=y1, keys=keys, seq=seq[0])

        roots = [mk(int(hx * f32(i)), int(hx * f32(i + 1)), 0, max_y - min_y, []) for i in range(n_ini)]
        for k, (x, y, s) in enumerate(cands):
            roots[int(f32(x) / hx)]["keys"].append(k)
        nodes = [r for r in roots if r["keys"]]

        def split(nd):
            hx_ = int(math.ceil(f32(nd["x1"] - nd["x0"]) / 2))
            hy_ = int(math.ceil(f32(nd["y1"] - nd["y0"]) / 2))
            mx, my = nd["x0"] + hx_, nd["y0"] + hy_
            q = [[], [], [], []]
            for k in nd["keys"]:
                x, y, _ = cands[k]
                if x < mx:
                    q[0 if y < my else 2].append(k)
                else:
                    q[1 if y < my else 3].append(k)
            boxes = [(nd["x0"], mx, nd["y0"], my), (mx, nd["x1"], nd["y0"], my), (nd["x0"], mx, my, nd["y1"]),
                     (mx, nd["x1"], my, nd["y1"])]
            return [mk(b[0], b[1], b[2], b[3], keys) for b, keys in zip(boxes, q) if keys]

        finish = False
        while not finish:
            prev = len(nodes)
            front, expandable, n_expand = [], [], 0
            rest = []
            for nd in nodes:
                if len(nd["keys"]) == 1:
                    rest.append(nd)
                    continue
                ch = split(nd)
                for c in ch:
                    front.insert(0, c)
                    if len(c["keys"]) > 1:
                        n_expand += 1
                        expandable.append(c)
            nodes = front + rest
            if len(nodes) >= n_target or len(nodes) == prev:
                finish = True
            elif len(nodes) + 3 * n_expand > n_target:
                while not finish:
                    prev2 = len(nodes)
                    order = sorted(expandable, key=lambda c: (len(c["keys"]), c["seq"]))
                    expandable = []
                    for nd in reversed(order):
                        ch = split(nd)
                        for c in ch:
                            nodes.insert(0, c)
                            if len(c["keys"]) > 1:
                                expandable.append(c)
                        nodes = [x for x in nodes if x is not nd]
                        if len(nodes) >= n_target:
                            break
                    if len(nodes) >= n_target or len(nodes) == prev2:
                        finish = True
        kept = []
        for nd in nodes:
            best = nd["keys"][0]
            for k in nd["keys"][1:]:
                if cands[k][2] > cands[best][2]:
                    best = k
            kept.append(best)
        return kept

    def ic_angle(self, im, x, y):
        m01 = m10 = 0
        for u in range(-15, 16):
            m10 += u * int(im[y, x + u])
        for v in range(1, 16):
            d = self.umax[v]
            vs = 0
            for u in range(-d, d + 1):
                p, m = int(im[y + v, x + u]), int(im[y - v, x + u])
                vs += p - m
                m10 += u * (p + m)
            m01 += v * vs
        return cv2.fastAtan2(float(m01), float(m10))

    def descriptor(self, blurred, x, y, angle):
        ang = f32(angle) * f32(np.pi / 180.0)
        a, b = f32(math.cos(float(ang))), f32(math.sin(float(ang)))  # correctly rounded (deviation 3)
        px = self.pattern[:, 0].astype(np.float32)
        py = self.pattern[:, 1].astype(np.float32)
        rr = np.rint(px * b + py * a).astype(np.int64)
        cc = np.rint(px * a - py * b).astype(np.int64)
        vals = blurred[y + rr, x + cc].astype(np.int32)
        bits = (vals[0::2] < vals[1::2]).astype(np.uint8)
        return np.packbits(bits, bitorder="little")

    def extract(self, img):
        levels = self.pyramid(img)
        out_kp, out_desc, stage = [], [], {}
        for lv, im in enumerate(levels):
            cands = self.cell_candidates(im)
            kept = self.quadtree(cands, 16, im.shape[1] - 16, 16, im.shape[0] - 16, self.quota[lv])
            stage["cand_%d" % lv] = np.array(cands, np.int32).reshape(-1, 3)
            stage["kept_%d" % lv] = np.array(kept, np.int32)
            if not kept:
                continue
            blurred = cv2.GaussianBlur(im, (7, 7), 2, 2, borderType=cv2.BORDER_REFLECT_101)
            for k in kept:
                x, y, s = cands[k]
                x, y = x + 16, y + 16
                ang = self.ic_angle(im, x, y)
                out_desc.append(self.descriptor(blurred, x, y, ang))
                sc = self.scale[lv]
                fx, fy = (f32(x), f32(y)) if lv == 0 else (f32(x) * sc, f32(y) * sc)
                out_kp.append((fx, fy, f32(int(f32(31) * sc)), f32(ang), f32(s), lv))
        kp = np.array(out_kp, np.float32).reshape(-1, 6)
        desc = np.array(out_desc, np.uint8).reshape(-1, 32)
        return kp, desc, stage, levels


def main():
    assert cv2.__version__.startswith("4.13"), cv2.__version__
    cases = {
        "small_a": (synth.synth_image(320, 240, seed=11, n_rect=120), (300, 1.2, 8, 20, 7)),
        "small_b": (synth.synth_image(400, 200, seed=12, n_rect=150), (500, 1.2, 6, 20, 7)),
        "small_noise": (synth.noise_image(200, 160, seed=13), (400, 1.2, 5, 20, 7)),
    }
    for name, (img, params) in cases.items():
        ex = PyExtractor(*params)
        kp, desc, stage, levels = ex.extract(img)
        path = os.path.join(HERE, "extract_%s.npz" % name)
        np.savez_compressed(path, image=img, params=np.array(params, np.float64), keypoints=kp, descriptors=desc,
                            level_last=levels[-1], **stage)
        print(name, "keypoints", len(kp), "->", os.path.relpath(path, ROOT))
    # full-size fixtures on BASELINE.json's three shapes (the image is regenerated from its seed by the tests, so only
    # its sha256 is stored): keypoints, descriptors, candidate counts and a sha256 per pyramid level + per level's
    # GaussianBlur (= a cv2 known answer for each of the 7 resize transitions and 8 blurs of the shape)
    import hashlib
    for shape in ("tum", "kitti", "euroc"):
        w, h = synth.SHAPES[shape]
        st = synth.SETTINGS[shape]
        params = (st["nfeatures"], st["scale_factor"], st["nlevels"], st["ini_th"], st["min_th"])
        img = synth.synth_image(w, h, seed=0)
        ex = PyExtractor(*params)
        kp, desc, stage, levels = ex.extract(img)
        sha = lambda a: np.frombuffer(hashlib.sha256(np.ascontiguousarray(a).tobytes()).digest(), np.uint8)
        path = os.path.join(HERE, "extract_full_%s.npz" % shape)
        np.savez_compressed(
            path, image_sha256=sha(img), shape=np.array([w, h], np.int32), seed=np.int32(0),
            params=np.array(params, np.float64), keypoints=kp, descriptors=desc, level_last=levels[-1],
            level_sha256=np.stack([sha(l) for l in levels]),
            blur_sha256=np.stack([sha(cv2.GaussianBlur(l, (7, 7), 2, 2, borderType=cv2.BORDER_REFLECT_101)) for l in levels]),
            level_dims=np.array([[l.shape[1], l.shape[0]] for l in levels], np.int32),
            cand_counts=np.array([len(stage["cand_%d" % lv]) for lv in range(len(levels))], np.int32),
            cand_sha256=np.stack([sha(stage["cand_%d" % lv].astype(np.int32)) for lv in range(len(levels))]),
            kept_counts=np.array([len(stage["kept_%d" % lv]) for lv in range(len(levels))], np.int32),
            cv2_version=np.array(cv2.__version__))
        print("full", shape, "keypoints", len(kp), "->", os.path.relpath(path, ROOT))
    # primitive known-answer vectors
    rng = np.random.default_rng(7)
    src = rng.integers(0, 256, (97, 131), dtype=np.uint8)
    yy = rng.integers(-70000, 70000, 512).astype(np.float32)
    xx = rng.integers(-70000, 70000, 512).astype(np.float32)
    at = np.array([cv2.fastAtan2(float(y), float(x)) for y, x in zip(yy, xx)], np.float32)
    det = cv2.FastFeatureDetector_create(20, True, cv2.FAST_FEATURE_DETECTOR_TYPE_9_16)
    roi = cv2.GaussianBlur(rng.integers(0, 256, (40, 37), dtype=np.uint8), (3, 3), 1)
    fk = np.array([(int(k.pt[0]), int(k.pt[1]), int(k.response)) for k in det.detect(roi)], np.int32)
    np.savez_compressed(os.path.join(HERE, "primitives.npz"), src=src,
                        resized=cv2.resize(src, (109, 81), interpolation=cv2.INTER_LINEAR),
                        blurred=cv2.GaussianBlur(src, (7, 7), 2, 2, borderType=cv2.BORDER_REFLECT_101),
                        atan_y=yy, atan_x=xx, atan=at, fast_roi=roi, fast_kps=fk)
    print("primitives ->", "tests/golden/primitives.npz")


if __name__ == "__main__":
    main()
