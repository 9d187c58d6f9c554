"""CPU: the ORB-SLAM3 guided-matching oracle (oracle/orb_guided.c, a C restatement of the reference's
ORBmatcher::SearchForInitialization and Frame::ComputeStereoMatches) and the NumPy restatement of the CUDA ray caster.
The reference holds no vectors for these functions and ORB-SLAM3 cannot be compiled here, so the oracle is checked for
what the reference lines promise: known-answer popcounts, one-to-one matches, thresholds, and stereo disparities that
agree with the geometry the images were rendered from."""
import numpy as np

from forest_slam_b200 import synth
from forest_slam_b200.api import ForestCamera
from oracle import render_numpy as rn
from tests import util


def test_descriptor_distance_known_answers():
    lib = util.orb_oracle()
    rng = np.random.default_rng(0)
    for _ in range(50):
        a, b = rng.integers(0, 256, 32, dtype=np.uint8), rng.integers(0, 256, 32, dtype=np.uint8)
        assert lib.orb_descriptor_distance(util.vp(a), util.vp(b)) == int(np.unpackbits(a ^ b).sum())


def _pair(seed, n=900):
    world = synth.make_world(n_trunks=80, n_ground=2500, n_canopy=2500)
    fr = synth.make_sequence(3, n, seed=synth.SEED0 + seed, world=world)
    ang = [synth.keypoint_angles(world.points.shape[0], f.landmark, 100 * seed + i) for i, f in enumerate(fr)]
    return fr, ang


def test_search_for_initialization_oracle_properties():
    fr, ang = _pair(1)
    a, b = fr[0], fr[2]
    n = a.kpts.shape[0]
    octv = np.zeros(n, np.int32)
    m12, nm, prev = util.orb_search_init_oracle(a.kpts, ang[0], octv, a.desc_bin, b.kpts, ang[2], octv, b.desc_bin, a.kpts)
    ok = m12 >= 0
    assert nm == int(ok.sum()) and nm > 300
    assert np.unique(m12[ok]).size == nm                                 # one-to-one (ORBmatcher.cc:710-716)
    right = a.landmark[ok] == b.landmark[m12[ok]]
    assert right.mean() > 0.97                                           # TH_LOW + ratio leave the true matches
    d = np.unpackbits(a.desc_bin[ok] ^ b.desc_bin[m12[ok]], axis=1).sum(1)
    assert d.max() <= 50                                                 # bestDist <= TH_LOW
    w = np.abs(b.kpts[m12[ok]] - a.kpts[ok])
    assert w.max() < 100                                                 # inside the search window
    assert np.array_equal(prev[ok], b.kpts[m12[ok]]) and np.array_equal(prev[~ok], a.kpts[~ok])
    # queries on a pyramid level > 0 are skipped (ORBmatcher.cc:666)
    o1 = octv.copy(); o1[::3] = 1
    m12b, _, _ = util.orb_search_init_oracle(a.kpts, ang[0], o1, a.desc_bin, b.kpts, ang[2], octv, b.desc_bin, a.kpts)
    assert np.all(m12b[::3] == -1)


def test_renderer_numpy_and_stereo_oracle_agree_with_geometry():
    cam = ForestCamera()
    world, T_l, T_r, fl, fr = util.stereo_scene(n_kpts=800, seed=2, n_trunks=40)
    imL, dispL = rn.render_forest(T_l, cam.fx, cam.fy, cam.cx, cam.cy, cam.baseline, world.ground_y, world.trunks)
    imR, _ = rn.render_forest(T_r, cam.fx, cam.fy, cam.cx, cam.cy, cam.baseline, world.ground_y, world.trunks)
    assert imL.shape == (600, 960) and imL.dtype == np.uint8 and imL.std() > 10
    # the right image is the left one displaced by the rendered disparity (textured surfaces, away from occlusions)
    ys, xs = np.mgrid[300:560:7, 200:900:11]
    d = dispL[ys, xs]
    xr = np.rint(xs - d).astype(int)
    ok = (d > 0) & (xr >= 0)
    assert (np.abs(imL[ys, xs][ok].astype(int) - imR[ys, xr][ok].astype(int)) <= 40).mean() > 0.6
    octv = np.zeros(fl.kpts.shape[0], np.int32)
    ur, dp, sad = util.orb_stereo_oracle(fl.kpts, octv, fl.desc_bin, fr.kpts, octv, fr.desc_bin, imL, imR,
                                         cam.baseline, cam.fx * cam.baseline)
    m = ur >= 0
    assert m.sum() > 100
    # matched keypoints: u_right agrees with the rendered disparity at the keypoint (where the landmark is the visible
    # surface) for most of them
    x, y = fl.kpts[m, 0], fl.kpts[m, 1]
    dd = dispL[np.rint(y).astype(int), np.rint(x).astype(int)]
    err = np.abs((x - ur[m]) - dd)
    assert np.median(err) < 1.0
    assert np.all(dp[m] > 0) and np.all((sad[m] >= 0))
