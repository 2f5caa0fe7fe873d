import sys, os
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from my_orb_slam2_b200 import orbfe
from oracle import pyoracle as po
g = np.load("tests/golden/extract_small_a.npz")
img = g["image"]; p = g["params"]; params = (int(p[0]), float(p[1]), int(p[2]), int(p[3]), int(p[4]))
print(img[16:24, 16:28])
roi = img[16:16+41, 16:16+38]
print("oracle thr20", po.fast9_16(np.ascontiguousarray(roi), 20)); print("oracle thr7", po.fast9_16(np.ascontiguousarray(roi), 7))
ex = orbfe.Extractor(*params, img.shape[1], img.shape[0]); kp, desc = ex.extract(img)
