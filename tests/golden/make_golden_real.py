"""Generates tests/golden/ambush_5_bm.npz: the reference's own tutorial stereo pair
(modules/ximgproc/tutorials/images/ambush_5_left.jpg / ambush_5_right.jpg, 1024x436) taken through the sample's default
path (samples/disparity_filtering.cpp:128-190: views halved, StereoBM(80, 7) left + right matcher, filter at full
resolution with lambda 8000 / sigma 1.5) with the REFERENCE filter itself (oracle/_ref, `make -C oracle ref`, 4 threads).

    python tests/golden/make_golden_real.py          (build container only: needs /root/reference and cv2)

Stored: the two JPEG files as byte arrays (the guide must be the very image the reference saw), the CV_16S
disparity maps cv2.StereoBM produced here (so the stored outputs stay valid whatever cv2 build decodes / matches later),
and the reference's filtered map, confidence map and ROI."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import ref as R  # noqa: E402
from tests.stereo_inputs import sample_matching  # noqa: E402

HERE = os.path.dirname(os.path.abspath(__file__))
IMAGES = "/root/reference/modules/ximgproc/tutorials/images"


def main():
    import cv2
    assert R.available(), "build oracle/_ref first: make -C oracle ref"
    R.set_num_threads(4)
    lj = np.fromfile(os.path.join(IMAGES, "ambush_5_left.jpg"), np.uint8)
    rj = np.fromfile(os.path.join(IMAGES, "ambush_5_right.jpg"), np.uint8)
    left, right = cv2.imdecode(lj, cv2.IMREAD_COLOR), cv2.imdecode(rj, cv2.IMREAD_COLOR)
    wls, dl, dr = sample_matching(left, right, "bm", downscale=True, max_disp=160)
    lo, ro, to, bo = wls._offsets
    dh, dw = dl.shape
    roi = (lo, to, dw - lo - ro, dh - to - bo)                                   # disparity_filters.cpp:228-233
    r = R.wls_filter(dl, left, dr, roi, True, 8000.0, 1.5, wls.getLRCthresh(), wls.getDepthDiscontinuityRadius())
    assert r["rc"] == 0 and r["roi"] == roi
    np.savez_compressed(os.path.join(HERE, "ambush_5_bm.npz"), left_jpg=lj, right_jpg=rj, left_sum=np.int64(left.astype(np.int64).sum()),
                        dl=dl, dr=dr, out=r["out"], conf=r["conf"], roi=np.array(roi), radius=np.int32(wls.getDepthDiscontinuityRadius()),
                        offsets=np.array(wls._offsets))
    print("ambush_5_bm.npz written:", os.path.getsize(os.path.join(HERE, "ambush_5_bm.npz")), "bytes; ROI", roi, "valid share",
          float((dl > 0).mean()))


if __name__ == "__main__":
    main()
