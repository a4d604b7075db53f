"""Seeded synthetic JRDB-shaped LiDAR frames: the reference ships no data, so
tests and the benchmark draw frames whose SHAPES follow its dataloader
(/root/reference/dataloader.py:187-195 boxes as cx,cy,cz,dx,dy,dz,heading; points x,y,z,intensity,
Multimodal_cfg/mil3.yaml:40-46) and whose statistics exercise every branch of the pooling path:
empty persons, persons with fewer than S points (wrap-duplicate), more than S (truncation),
overlapping boxes, ground-plane clutter.
"""
import numpy as np

POINT_CLOUD_RANGE = (-25.0, -25.0, -1.5, 25.0, 25.0, 2.5)


def make_boxes(rng, M, overlap_frac=0.1):
    boxes = np.zeros((M, 7), dtype=np.float32)
    centres = []
    for m in range(M):
        for _ in range(64):
            r, a = rng.uniform(1.5, 15.0), rng.uniform(-np.pi, np.pi)
            c = np.array([r * np.cos(a), r * np.sin(a)])
            if all(np.hypot(*(c - o)) >= 0.8 for o in centres):
                break
        centres.append(c)
        dz = rng.uniform(1.4, 1.95)
        boxes[m] = [c[0], c[1], -0.5 + dz / 2 + rng.normal(0, 0.05), rng.uniform(0.4, 0.9), rng.uniform(0.4, 0.9), dz,
                    rng.uniform(-np.pi, np.pi)]
    n_pairs = int(M * overlap_frac / 2)          # deliberately overlapping pairs
    for p in range(n_pairs):
        a, b = 2 * p, 2 * p + 1
        if b < M:
            boxes[b, 0:2] = boxes[a, 0:2] + rng.uniform(-0.3, 0.3, 2).astype(np.float32)
    return boxes


def make_frame(seed, N=65536, M=40, density_scale=1.0, empty_frac=0.05):
    """-> points (N,4) float32 [x,y,z,intensity], boxes (M,7) float32."""
    rng = np.random.default_rng(seed)
    boxes = make_boxes(rng, M)
    fg = []
    empty = rng.random(M) < empty_frac
    if M >= 8 and empty_frac > 0 and not empty.any():
        empty[int(rng.integers(M))] = True       # the empty-flag path is always exercised
    for m in range(M):
        n_p = 0 if empty[m] else int(np.clip(rng.lognormal(np.log(150.0 * density_scale), 0.8), 0, 2000 * density_scale))
        if n_p == 0:
            continue
        cx, cy, cz, dx, dy, dz, rz = boxes[m]
        loc = rng.uniform(-0.5, 0.5, (n_p, 3)) * np.array([dx, dy, dz])
        c, s = np.cos(rz), np.sin(rz)
        x = loc[:, 0] * c - loc[:, 1] * s + cx
        y = loc[:, 0] * s + loc[:, 1] * c + cy
        fg.append(np.stack([x, y, loc[:, 2] + cz], axis=1))
    fg = np.concatenate(fg, axis=0) if fg else np.zeros((0, 3))
    fg = fg[:N]
    n_bg = N - fg.shape[0]
    n_ground = int(n_bg * 0.6)
    rad, ang = 25.0 * np.sqrt(rng.random(n_ground)), rng.uniform(-np.pi, np.pi, n_ground)
    ground = np.stack([rad * np.cos(ang), rad * np.sin(ang), -0.5 + rng.normal(0, 0.02, n_ground)], axis=1)
    rest = np.stack([rng.uniform(-25, 25, n_bg - n_ground), rng.uniform(-25, 25, n_bg - n_ground),
                     rng.uniform(-1, 2, n_bg - n_ground)], axis=1)
    xyz = np.concatenate([fg, ground, rest], axis=0)
    lo, hi = np.array(POINT_CLOUD_RANGE[:3]), np.array(POINT_CLOUD_RANGE[3:])
    xyz = np.clip(xyz, lo + 1e-3, hi - 1e-3)
    for m in np.nonzero(empty)[0]:               # lift clutter out of the persons meant to be empty
        cx, cy, cz, dx, dy, dz, rz = boxes[m]
        near = (np.abs(xyz[:, 0] - cx) < dx + dy) & (np.abs(xyz[:, 1] - cy) < dx + dy) & \
               (np.abs(xyz[:, 2] - cz) < dz / 2 + 0.05)
        xyz[near, 2] = cz + dz / 2 + 0.2
    pts = np.concatenate([xyz, rng.random((N, 1))], axis=1).astype(np.float32)
    rng.shuffle(pts, axis=0)   # DataProcessor shuffle_points (mil3.yaml:51-54)
    return pts, boxes


def make_batch(seed, B, N=65536, M=40, **kw):
    frames = [make_frame(seed * 1000 + b, N, M, **kw) for b in range(B)]
    return np.stack([f[0] for f in frames]), np.stack([f[1] for f in frames])
