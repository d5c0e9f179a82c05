"""Row N2 timing on the GPU box: candidate generation + kernel 2 per image, host enumeration
(itertools + set(), the reference's way) against the device path, on the C4 star fields."""
import json, os, sys, time
import cv2
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import ssp_b200
from ssp_b200 import synthetic as synth

W, H, F = 1342, 894, 1135.0
n_img = int(sys.argv[1]) if len(sys.argv) > 1 else 8
xy, br, _ = synth.star_fields(n_img, 2000, seed=2, img_w=W, img_h=H, focal=F)
frame = np.zeros((H, W, 3), np.uint8)
eng = ssp_b200.Engine(0)
out = {}
for mode in ("host", "device", "host", "device"):
    bf = ssp_b200.BFPolygonMatcher("/tmp", "a", "b", 0, 1, frame.shape, F, engine=eng, write_debug_images=False,
                                   candidate_generation=mode)
    sess = bf._session({})
    t_total, n_c, n_p = 0.0, [], []
    for i in range(n_img):
        kps = tuple(cv2.KeyPoint(float(p[0]), float(p[1]), 4.0) for p in xy[i].astype(np.float32))
        bf._brightness_cache[(id(frame), len(kps))] = br[i]          # brightness is an input of this workload
        t0 = time.perf_counter()
        if mode == "device":
            cands, polys = bf._measure_with_device_candidates(sess, i, kps, frame)
        else:
            cands = bf.reduce_amount_of_keypoints_and_create_polygon_spanning_tuples_of_keypoints(kps, f"i{i}", frame, 5, 15000, _brightness=br[i])
            polys = bf.get_polygons_characteristics(kps, frame, "t", 5, cands, _img_id=i, _session=sess)
        t_total += time.perf_counter() - t0
        n_c.append(len({tuple(sorted(c)) for c in cands}))
        n_p.append(len(polys["keypoint_indices_rotated_list"]))
    out[mode] = {"ms_per_image": t_total / n_img * 1e3, "unique_candidates": n_c, "polygons": n_p}
assert out["host"]["unique_candidates"] == out["device"]["unique_candidates"] and out["host"]["polygons"] == out["device"]["polygons"]
print(json.dumps({"what": "candidate generation (regions, Ward clustering on the host in both modes, per-cluster selection, C(k,5) "
                          "enumeration + de-duplication) + kernel 2 + bank fetch, per image, C4 star fields (2000 stars)",
                  "images": n_img, **out}))
