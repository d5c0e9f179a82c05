"""Synthetic workloads of the shapes BASELINE.json names.

  descriptor_bank      C3 / C5: images x 256-bit ORB-like descriptors with planted matches
  tie_stress_bank      adversarial variant: tiny alphabet + exact duplicates -> many equal distances
  star_fields          C4: pinhole views of one random sky, ~n stars per view
  render_star_image    uint8 BGR frame with the stars drawn as discs (for the brightness path)
  polygon_feature_bank normalised 15-dim polygon rows for kernel-3 stress runs
"""
import numpy as np

C3 = dict(n_images=21, n_desc=50_000, seed=0)     # "synthetic 21 images x 50k ORB descriptors"
C5 = dict(n_images=200, n_desc=20_000, seed=1)    # "synthetic 200-image sweep x 20k descriptors"
C4 = dict(n_images=21, n_stars=2_000, seed=2, img_w=1342, img_h=894, focal=1135.0)


def descriptor_bank(n_images, n_desc, seed=0, world=200_000, stride_frac=0.6, flip_p=0.05):
    """A 'world' of random 256-bit codes; image k copies a shuffled window of the world (adjacent
    images share 1 - stride_frac = 40 % of their codes) and flips every bit with probability
    flip_p, so true matches sit near distance 256*2p(1-p) ~ 24 and non-matches near 128, like ORB.
    Uniform random rows would leave no ratio-test survivors."""
    rng = np.random.default_rng(seed)
    world = max(int(world), int(n_desc))
    codes = rng.integers(0, 256, size=(world, 32), dtype=np.uint8)
    stride = int(n_desc * stride_frac)
    out = []
    for k in range(n_images):
        ids = (k * stride + rng.permutation(n_desc)) % world
        rows = codes[ids]
        flips = np.packbits(rng.random((n_desc, 256), dtype=np.float32) < flip_p, axis=1)
        out.append(np.ascontiguousarray(rows ^ flips))
    return out


def tie_stress_bank(n_images, n_desc, seed=0, dup_frac=0.2):
    """Rows whose bytes come from a 4-value alphabet (distances cluster on few integers) plus
    exact duplicates inside and across images: ties are the norm, which is what the
    lowest-index tie-break has to survive."""
    rng = np.random.default_rng(seed)
    alphabet = np.array([0x00, 0x0F, 0xF0, 0xFF], dtype=np.uint8)
    pool = alphabet[rng.integers(0, 4, size=(max(8, n_desc // 4), 32))]
    out = []
    for _ in range(n_images):
        rows = alphabet[rng.integers(0, 4, size=(n_desc, 32))]
        n_dup = int(n_desc * dup_frac)
        if n_dup:
            rows[rng.integers(0, n_desc, size=n_dup)] = pool[rng.integers(0, len(pool), size=n_dup)]
        out.append(np.ascontiguousarray(rows))
    return out


def all_pairs(n_images):
    return [(i, j) for i in range(n_images) for j in range(i + 1, n_images)]


# ---- star fields ---------------------------------------------------------------------------

def _rot_yaw_pitch(yaw_deg, pitch_deg):
    y, p = np.deg2rad(yaw_deg), np.deg2rad(pitch_deg)
    ry = np.array([[np.cos(y), 0, np.sin(y)], [0, 1, 0], [-np.sin(y), 0, np.cos(y)]])
    rp = np.array([[1, 0, 0], [0, np.cos(p), -np.sin(p)], [0, np.sin(p), np.cos(p)]])
    return ry @ rp


def camera_poses(n_images):
    """8 + 8 + 4 + 1 all-sky plan (like the reference's shot lists): two rings of 8, one of 4,
    zenith; repeated with a yaw offset when more views are asked for."""
    base = [(45.0 * k, 10.0) for k in range(8)] + [(45.0 * k + 22.5, 38.0) for k in range(8)] + \
           [(90.0 * k, 64.0) for k in range(4)] + [(0.0, 90.0)]
    poses = []
    rep = 0
    while len(poses) < n_images:
        poses += [(y + 7.0 * rep, p) for y, p in base]
        rep += 1
    return poses[:n_images]


def star_fields(n_images=21, n_stars=2000, seed=2, img_w=1342, img_h=894, focal=1135.0,
                integer_pixels=True):
    """Returns (xy_list, brightness_list, sky_id_list): per view the pixel coordinates (float64),
    brightness in [60, 255] and the sky index of every visible star."""
    rng = np.random.default_rng(seed)
    # solid angle of the pinhole frustum -> sky density giving ~n_stars per view
    hx, hy = np.arctan(img_w / 2 / focal), np.arctan(img_h / 2 / focal)
    omega = 4.0 * np.arcsin(np.sin(hx) * np.sin(hy))
    n_sky = int(n_stars * 4 * np.pi / omega)
    v = rng.normal(size=(n_sky, 3))
    v /= np.linalg.norm(v, axis=1, keepdims=True)
    bright = rng.uniform(60, 255, size=n_sky)
    xy_l, b_l, id_l = [], [], []
    for yaw, pitch in camera_poses(n_images):
        cam = v @ _rot_yaw_pitch(yaw, -pitch)  # camera looks along +z
        front = cam[:, 2] > 1e-6
        x = focal * cam[:, 0] / np.where(front, cam[:, 2], 1.0) + img_w // 2
        y = focal * cam[:, 1] / np.where(front, cam[:, 2], 1.0) + img_h // 2
        if integer_pixels:
            x, y = np.rint(x), np.rint(y)
        vis = front & (x >= 3) & (x < img_w - 3) & (y >= 3) & (y < img_h - 3)
        ids = np.nonzero(vis)[0]
        xy = np.stack([x[ids], y[ids]], axis=1).astype(np.float64)
        if integer_pixels and len(xy):
            _, first = np.unique(xy, axis=0, return_index=True)  # one star per pixel
            keep = np.sort(first)
            xy, ids = xy[keep], ids[keep]
        xy_l.append(xy)
        b_l.append(bright[ids].copy())
        id_l.append(ids)
    return xy_l, b_l, id_l


def render_star_image(xy, brightness, img_w=1342, img_h=894, radius=2):
    """uint8 (h, w, 3) BGR frame: black sky, every star a filled disc of its brightness."""
    import cv2
    img = np.zeros((img_h, img_w, 3), dtype=np.uint8)
    for (x, y), b in zip(xy, brightness):
        v = int(b)
        cv2.circle(img, (int(x), int(y)), radius, (v, v, v), thickness=-1)
    return img


def random_candidates(n_stars, n_candidates, seed=0):
    """n_candidates random 5-subsets of range(n_stars) (int32 (n, 5))."""
    rng = np.random.default_rng(seed)
    out = np.empty((n_candidates, 5), dtype=np.int32)
    for i in range(n_candidates):
        out[i] = rng.choice(n_stars, size=5, replace=False)
    return out


def polygon_feature_bank(n_images, n_polygons, seed=3):
    """Random but plausibly normalised 15-dim rows: angles/pi in (0.1, 0.9), sides and
    brightness scaled so each row's maximum is 1."""
    rng = np.random.default_rng(seed)
    out = []
    for _ in range(n_images):
        a = rng.uniform(0.1, 0.9, size=(n_polygons, 5))
        s = rng.uniform(0.05, 1.0, size=(n_polygons, 5))
        b = rng.uniform(0.2, 1.0, size=(n_polygons, 5))
        s /= s.max(axis=1, keepdims=True)
        b /= b.max(axis=1, keepdims=True)
        out.append(np.ascontiguousarray(np.concatenate([a, s, b], axis=1)))
    return out
