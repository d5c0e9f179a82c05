"""Mirror of the two functions of the reference's image_processors.py that sit directly in front of the
star-polygon path: `get_star_brightness` (image_processors.py:222-254) and `compute_star_features`
(image_processors.py:43-219).

`compute_star_features` is the reference's star finder: Canny edges -> contours -> one keypoint per
contour (centroid, contour area) -> keep the `n_brightest_stars` by `get_star_brightness` -> ORB
descriptors at those keypoints.  Its time is the brightness loop -- a full-frame mask, bitwise_and and
np.sort PER STAR (90 % of the polygon path's wall time, 148 s of 164 s per pair) -- which is
exactly what `ssp_star_brightness` (csrc/brightness.cu, row N1) computes for all stars of a frame in
one launch, integer-exact.  Edge detection, contour tracing and ORB description stay on the
reference's OpenCV calls, so the returned keypoints and descriptors are the reference's, bit for bit
(tests/test_polygon_gpu.py: the recorded output of the unchanged reference on the eisenberg frames).
This covers the brightness half of SURVEY §8(f) N4; a CUDA Canny / contour tracer is not built.
"""
import os

import cv2 as cv
import numpy as np

from .custom_bf_polygon_matcher import _debug_images_enabled, disc_offsets
from .engine import Engine

BRIGHTNESS_RADIUS = 10   # image_processors.py:246
BRIGHTNESS_TOP_K = 10    # image_processors.py:253


def star_brightnesses(keypoints, workscale_image, engine: Engine):
    """get_star_brightness for every keypoint of one image in ONE kernel launch: int per keypoint."""
    gray = cv.cvtColor(workscale_image, cv.COLOR_BGR2GRAY)
    centers = np.array([[int(kp.pt[0]), int(kp.pt[1])] for kp in keypoints], dtype=np.int32).reshape(-1, 2)
    if len(centers) == 0:
        return np.zeros(0, dtype=np.int32)
    return engine.star_brightness(gray, centers, disc_offsets(BRIGHTNESS_RADIUS), BRIGHTNESS_TOP_K)


def get_star_brightness(keypoint, workscale_image, engine: Engine):
    """image_processors.py:222-254 for one keypoint: sum of the 10 largest grey values inside the filled
    circle of radius 10 around int(keypoint.pt)."""
    return int(star_brightnesses([keypoint], workscale_image, engine)[0])


def compute_star_features(img, img_name, output_dir, n_brightest_stars=500, engine: Engine = None,
                          write_debug_images=None):
    """image_processors.py:43-219.  Returns a cv2.detail.ImageFeatures whose keypoints are the detected
    stars (brightest first) with their ORB descriptors."""
    if engine is None:
        raise ValueError("compute_star_features needs an Engine: there is no CPU fallback")
    debug = _debug_images_enabled(write_debug_images)
    orig_image = img.copy()
    img_gray = cv.cvtColor(img, cv.COLOR_BGR2GRAY)
    edges = cv.Canny(img_gray, 201, 81)                                          # :88-92
    contours, _ = cv.findContours(edges, cv.RETR_TREE, cv.CHAIN_APPROX_NONE)    # :94-98
    if debug:
        os.makedirs(output_dir, exist_ok=True)
        cv.imwrite(os.path.join(output_dir, f"{img_name}___05_canny_edges.jpg"), edges)
        cv.imwrite(os.path.join(output_dir, f"{img_name}___06_contours.jpg"),
                   cv.drawContours(orig_image.copy(), contours, -1, (0, 255, 0), 2))
    keypoints, seen = [], set()
    for c in contours:                                                           # :110-135
        m = cv.moments(c)
        try:
            cx, cy = int(m["m10"] / m["m00"]), int(m["m01"] / m["m00"])
        except ZeroDivisionError:
            cx, cy = int(np.mean(c[:, 0, 0])), int(np.mean(c[:, 0, 1]))
            if c.shape[1] != 1:
                raise ValueError()
        if (cx, cy) in seen:                                                     # one keypoint per centre
            continue
        seen.add((cx, cy))
        kp = cv.KeyPoint()
        kp.angle = -1.0
        kp.class_id = -1
        kp.octave = 0
        kp.pt = (cx, cy)
        kp.response = 0.0
        kp.size = cv.contourArea(c)
        keypoints.append(kp)
    # :160-168 -- the reference's brightness loop, here one launch for the whole frame
    brightness = star_brightnesses(keypoints, orig_image, engine).tolist()
    ranked = sorted(zip(keypoints, brightness), key=lambda x: x[1])[::-1][:n_brightest_stars]
    brightest = [kp for kp, _ in ranked]
    orb = cv.ORB_create(nfeatures=len(keypoints) + 1, edgeThreshold=0)            # :154-157
    keypoints_recomputed, descriptors = orb.compute(orig_image, brightest)       # :170
    if debug:
        cv.imwrite(os.path.join(output_dir, f"{img_name}___07_orb_computed_{len(brightest)}_keypoints.jpg"),
                   cv.drawKeypoints(orig_image.copy(), keypoints_recomputed, None, color=(0, 255, 0), flags=0))
    feats = cv.detail.computeImageFeatures2(cv.ORB_create(), np.zeros(orig_image.shape, dtype=np.uint8))   # :189-193
    feats.descriptors = cv.UMat(descriptors)
    feats.img_idx = -1
    feats.img_size = orig_image.shape[:2][::-1]
    feats.keypoints = keypoints_recomputed
    return feats
