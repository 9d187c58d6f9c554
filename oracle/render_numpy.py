"""TEST INFRASTRUCTURE (oracle): NumPy restatement of forest_slam_b200/csrc/render.cu (the CUDA ray caster that renders
the synthetic stereo image pairs).  The reference has no renderer (its images come from a rosbag, stereo_slam.py:35);
this pins the CUDA kernel: same operations in the same order in float64, so the images are bit-identical."""
import numpy as np


def _hash3(x, y, z):
    x = x.astype(np.int64).astype(np.uint32)
    y = y.astype(np.int64).astype(np.uint32)
    z = z.astype(np.int64).astype(np.uint32)
    with np.errstate(over="ignore"):
        h = (x * np.uint32(0x8da6b343)) ^ (y * np.uint32(0xd8163841)) ^ (z * np.uint32(0xcb1ab31f))
        h ^= h >> np.uint32(13)
        h = h * np.uint32(0x5bd1e995)
        h ^= h >> np.uint32(15)
    return h


def _solid_texture(X, Y, Z):
    a = (_hash3(np.floor(X * 20.0), np.floor(Y * 20.0), np.floor(Z * 20.0)) & np.uint32(255)).astype(np.int64)
    b = (_hash3(np.floor(X * 2.5), np.floor(Y * 2.5), np.floor(Z * 2.5)) & np.uint32(255)).astype(np.int64)
    return (a * 3 + b * 5) >> 3


def render_forest(T_wc, fx, fy, cx, cy, baseline, ground_y, trunks, width=960, height=600):
    """-> (image [H, W] u8, disparity [H, W] f32); trunks [n, 4] = x, z, radius, height."""
    T_wc = np.asarray(T_wc, dtype=np.float64)
    R, o = T_wc[:3, :3], T_wc[:3, 3]
    ix, iy = np.meshgrid(np.arange(width, dtype=np.float64), np.arange(height, dtype=np.float64))
    dcx, dcy = (ix - cx) / fx, (iy - cy) / fy
    dx = R[0, 0] * dcx + R[0, 1] * dcy + R[0, 2]
    dy = R[1, 0] * dcx + R[1, 1] * dcy + R[1, 2]
    dz = R[2, 0] * dcx + R[2, 1] * dcy + R[2, 2]
    tbest = np.full((height, width), 1e30)
    kind = np.zeros((height, width), np.int32)
    with np.errstate(divide="ignore", invalid="ignore"):
        t = (ground_y - o[1]) / dy
    g = (dy > 1e-12) & (t > 0.0) & (t < tbest)
    tbest[g] = t[g]
    kind[g] = 1
    a = dx * dx + dz * dz
    for tr in np.asarray(trunks, dtype=np.float64).reshape(-1, 4):
        ox, oz, r = o[0] - tr[0], o[2] - tr[1], tr[2]
        b = 2.0 * (ox * dx + oz * dz)
        c = ox * ox + oz * oz - r * r
        disc = b * b - 4.0 * a * c
        ok = (disc >= 0.0) & (a >= 1e-18)
        with np.errstate(divide="ignore", invalid="ignore"):
            t = (-b - np.sqrt(np.where(ok, disc, 0.0))) / (2.0 * a)
        y = o[1] + t * dy
        ok &= (t > 0.0) & (t < tbest) & ~(y > ground_y) & ~(y < ground_y - tr[3])
        tbest[ok] = t[ok]
        kind[ok] = 2
    X, Y, Z = o[0] + tbest * dx, o[1] + tbest * dy, o[2] + tbest * dz
    hit = kind > 0
    tex = np.zeros((height, width), np.int64)
    tex[hit] = _solid_texture(X[hit], Y[hit], Z[hit])
    v = np.where(kind == 0, 200 + (iy.astype(np.int64) * 40) // height,
                 np.where(kind == 1, 30 + (tex * 5) // 8, 60 + (tex * 3) // 4))
    disp = np.zeros((height, width), np.float32)
    disp[hit] = ((fx * baseline) / tbest[hit]).astype(np.float32)
    return np.minimum(v, 255).astype(np.uint8), disp
