"""Stage-by-stage GPU vs oracle diff for one image (debug aid; run on the GPU box)."""
import sys, os
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from my_orb_slam2_b200 import orbfe, synth
from oracle import pyoracle as po

def main():
    name = sys.argv[1] if len(sys.argv) > 1 else "small_a"
    if name.startswith("small"):
        g = np.load("tests/golden/extract_%s.npz" % name)
        img = g["image"]; p = g["params"]; params = (int(p[0]), float(p[1]), int(p[2]), int(p[3]), int(p[4]))
    else:
        s = synth.SETTINGS[name]; w, h = synth.SHAPES[name]
        img = synth.synth_image(w, h, 0); params = (s["nfeatures"], s["scale_factor"], s["nlevels"], s["ini_th"], s["min_th"])
    oex = po.Extractor(*params); okp, odesc = oex.extract(img)
    ex = orbfe.Extractor(*params, img.shape[1], img.shape[0]); kp, desc = ex.extract(img)
    print("n gpu", len(kp), "n oracle", len(okp))
    for lv in range(params[2]):
        a, b = ex.level(lv), oex.level(lv)
        print("level", lv, a.shape, "pyr mismatches", int((a != b).sum()))
        xs, ys, sc = ex.candidates(lv); oxs, oys, osc = oex.candidates(lv)
        gs = set(zip(xs.tolist(), ys.tolist(), sc.tolist())); os_ = set(zip(oxs.tolist(), oys.tolist(), osc.tolist()))
        print("   cand gpu", len(xs), "oracle", len(oxs), "only gpu", len(gs - os_), "only oracle", len(os_ - gs),
              "order ok", len(xs) == len(oxs) and np.array_equal(xs, oxs) and np.array_equal(ys, oys))
        if gs - os_: print("     e.g. only gpu", sorted(gs - os_)[:5])
        if os_ - gs: print("     e.g. only oracle", sorted(os_ - gs)[:5])
        kx, ky, ks = ex.level_keypoints(lv); ok = oex.level_keypoints(lv)
        same = len(kx) == len(ok) and np.array_equal(kx + 16, ok["x"].astype(np.int32)) and np.array_equal(ky + 16, ok["y"].astype(np.int32))
        print("   kept gpu", len(kx), "oracle", len(ok), "quota", oex.info.features_per_level[lv], "same", same)
        if not same:
            gk = set(zip((kx + 16).tolist(), (ky + 16).tolist())); okk = set(zip(ok["x"].astype(int).tolist(), ok["y"].astype(int).tolist()))
            print("     set diff: only gpu", len(gk - okk), "only oracle", len(okk - gk))
    if len(kp) == len(okp):
        for f in ("x", "y", "size", "angle", "response", "octave"):
            print(f, "mismatch", int((kp[f] != okp[f]).sum()))
        print("desc rows mismatch", int((desc != odesc).any(1).sum()))

main()
