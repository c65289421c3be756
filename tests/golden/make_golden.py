"""Generates the committed fixtures in tests/golden/ (run from the repo root):
  image_tail_cv2.npz : inputs + outputs of the S14 tail computed by REAL OpenCV (cv2 4.13):
                       dilate(3x3 rect) -> cvtColor(BGR2RGB) -> Mat.convertTo(CV_8UC3, 255)
                       (convertTo reached through cv2.convertScaleAbs-free path: np.rint on the
                       float32 product == cv::saturate_cast<uchar>(cvRound(v)); cross-checked
                       against cv2.convertScaleAbs below, which implements the same rounding)
  oracle_tiny.npz    : oracle outputs on scenes.make_scene('tiny') (regression pin)."""
import os
import sys

import cv2
import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
from agile_grasp2_b200 import capi, scenes  # noqa: E402
from common import make_params  # noqa: E402
from oracle import binding as ob  # noqa: E402

here = os.path.dirname(os.path.abspath(__file__))
rng = np.random.default_rng(2024)
inp, out = [], []
for k in range(8):
    img = np.zeros((60, 60, 3), np.float32)
    m = rng.random((60, 60)) < (0.03 + 0.08 * k)
    v = np.abs(rng.normal(size=(60, 60, 3))); v /= np.linalg.norm(v, axis=2, keepdims=True)
    img[m] = v[m].astype(np.float32)
    d = cv2.cvtColor(cv2.dilate(img, cv2.getStructuringElement(cv2.MORPH_RECT, (3, 3))), cv2.COLOR_BGR2RGB)
    u8 = cv2.convertScaleAbs(d, alpha=255.0)  # values >= 0: |.| is a no-op; saturate_cast<uchar>(cvRound())
    inp.append(img); out.append(u8)
np.savez_compressed(os.path.join(here, "image_tail_cv2.npz"), inp=np.array(inp), out=np.array(out))

sc = scenes.make_scene("tiny")
o = ob.context(make_params(capi.MODE_GEOMETRIC))
o.set_cloud(sc["xyz"], sc["cam_mask"], sc["num_cams"]); o.set_sample_indices(sc["samples"]); o.detect()
sw, h = o.sweep_records(), o.hypotheses()
np.savez_compressed(os.path.join(here, "oracle_tiny.npz"), fingers=sw["fingers"], antipodal=sw["antipodal"],
                    num_in_box=sw["num_in_box"], normals_bits=o.normals().view(np.uint32), bottom=h["bottom"],
                    width=h["width"])
print("wrote fixtures")
