"""GPU: guided ORB matching (csrc/orb_guided.cu) and the CUDA ray caster (csrc/render.cu) through the C ABI against the
oracle (oracle/orb_guided.c = ORB_SLAM3/src/ORBmatcher.cc:648-762, src/Frame.cc:811-981 restated; oracle/render_numpy.py).
Integer / index outputs bit-exact; the float outputs (u_right, depth) are the reference's float32 arithmetic: bit-exact."""
import numpy as np
import pytest

from forest_slam_b200 import synth
from forest_slam_b200.api import ForestCamera
from oracle import render_numpy as rn
from tests import util

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def fsb():
    import torch
    if not torch.cuda.is_available():
        pytest.skip("needs a GPU")
    from forest_slam_b200 import api
    return api


def test_render_forest_bit_exact(fsb):
    import torch
    cam = ForestCamera()
    world, T_l, T_r, _, _ = util.stereo_scene(n_kpts=100, seed=5, n_trunks=50)
    tr = torch.from_numpy(world.trunks).cuda()
    for T in (T_l, T_r):
        img, disp = fsb.render_forest(T, tr, ground_y=world.ground_y)
        ref_img, ref_disp = rn.render_forest(T, cam.fx, cam.fy, cam.cx, cam.cy, cam.baseline, world.ground_y, world.trunks)
        assert np.array_equal(img.cpu().numpy(), ref_img)
        assert np.array_equal(disp.cpu().numpy(), ref_disp)


@pytest.mark.parametrize("check", [True, False])
def test_search_for_initialization_vs_oracle(fsb, check):
    import torch
    world = synth.make_world(n_trunks=80, n_ground=2500, n_canopy=2500)
    frames = synth.make_sequence(6, 1500, world=world)
    n_max = 1536
    F = len(frames)
    kp = np.zeros((F, n_max, 2), np.float32)
    ang = np.zeros((F, n_max), np.float32)
    octv = np.zeros((F, n_max), np.int32)
    desc = np.zeros((F, n_max, 32), np.uint8)
    counts = np.array([1500, 1500, 1100, 1500, 700, 1500], np.int32)       # ragged banks
    rng = np.random.default_rng(3)
    for f, fr in enumerate(frames):
        c = counts[f]
        kp[f, :c], desc[f, :c] = fr.kpts[:c], fr.desc_bin[:c]
        ang[f, :c] = synth.keypoint_angles(world.points.shape[0], fr.landmark[:c], 10 + f)
        octv[f, :c] = (rng.random(c) < 0.15).astype(np.int32) * rng.integers(1, 4, c)   # some keypoints on levels 1..3
    pa, pb = [0, 1, 2, 3, 0, 4], [1, 2, 3, 4, 5, 5]
    prev = np.stack([kp[a] for a in pa]).copy()
    out = fsb.orb_search_for_initialization(torch.from_numpy(kp).cuda(), torch.from_numpy(octv).cuda(),
                                            torch.from_numpy(desc).cuda(), torch.from_numpy(counts).cuda(), pa, pb,
                                            torch.from_numpy(prev).cuda(), angle=torch.from_numpy(ang).cuda(),
                                            window_size=100, nn_ratio=0.9, check_orientation=check)
    m12 = out["matches12"].cpu().numpy()
    pm = out["prev_matched"].cpu().numpy()
    nm = out["n_matches"].cpu().numpy()
    total = 0
    for p, (a, b) in enumerate(zip(pa, pb)):
        ca, cb = counts[a], counts[b]
        ref, n_ref, prev_ref = util.orb_search_init_oracle(kp[a, :ca], ang[a, :ca], octv[a, :ca], desc[a, :ca], kp[b, :cb],
                                                           ang[b, :cb], octv[b, :cb], desc[b, :cb], prev[p, :ca],
                                                           check=check)
        assert np.array_equal(m12[p, :ca], ref), p
        assert nm[p] == n_ref
        assert np.array_equal(pm[p, :ca], prev_ref)
        total += n_ref
    assert total > 1500
    # a second call continues from the updated search centres (vbPrevMatched is in/out): still equal to the oracle
    out2 = fsb.orb_search_for_initialization(torch.from_numpy(kp).cuda(), torch.from_numpy(octv).cuda(),
                                             torch.from_numpy(desc).cuda(), torch.from_numpy(counts).cuda(), pa, pb,
                                             out["prev_matched"].clone(), angle=torch.from_numpy(ang).cuda(),
                                             window_size=30, nn_ratio=0.9, check_orientation=check)
    for p, (a, b) in enumerate(zip(pa, pb)):
        ca, cb = counts[a], counts[b]
        ref, n_ref, _ = util.orb_search_init_oracle(kp[a, :ca], ang[a, :ca], octv[a, :ca], desc[a, :ca], kp[b, :cb],
                                                    ang[b, :cb], octv[b, :cb], desc[b, :cb], pm[p, :ca], window=30,
                                                    check=check)
        assert np.array_equal(out2["matches12"][p, :ca].cpu().numpy(), ref), p


def test_stereo_matches_vs_oracle_and_feeds_backprojection(fsb):
    import torch
    cam = ForestCamera()
    scenes = [util.stereo_scene(n_kpts=1200, seed=s, n_trunks=50, pose_index=2 + s) for s in range(3)]
    F, n_max = len(scenes), 1280
    kpL = np.zeros((F, n_max, 2), np.float32); kpR = np.zeros_like(kpL)
    dL = np.zeros((F, n_max, 32), np.uint8); dR = np.zeros_like(dL)
    oL = np.zeros((F, n_max), np.int32); oR = np.zeros_like(oL)
    cL = np.array([1200, 1000, 1200], np.int32); cR = np.array([1200, 1200, 900], np.int32)
    imL = torch.empty((F, 600, 960), dtype=torch.uint8, device="cuda"); imR = torch.empty_like(imL)
    for f, (world, T_l, T_r, fl, fr) in enumerate(scenes):
        kpL[f, :cL[f]], dL[f, :cL[f]] = fl.kpts[:cL[f]], fl.desc_bin[:cL[f]]
        kpR[f, :cR[f]], dR[f, :cR[f]] = fr.kpts[:cR[f]], fr.desc_bin[:cR[f]]
        oR[f, :cR[f]:17] = 2    # a few right keypoints two levels up: excluded by the octave gate (Frame.cc:876)
        tr = torch.from_numpy(world.trunks).cuda()
        imL[f] = fsb.render_forest(T_l, tr, ground_y=world.ground_y, want_disparity=False)[0]
        imR[f] = fsb.render_forest(T_r, tr, ground_y=world.ground_y, want_disparity=False)[0]
    t = lambda a: torch.from_numpy(a).cuda()
    out = fsb.orb_stereo_matches(t(kpL), t(oL), t(dL), t(cL), t(kpR), t(oR), t(dR), t(cR), imL, imR)
    hL, hR = imL.cpu().numpy(), imR.cpu().numpy()
    n_match = 0
    for f in range(F):
        ur, dp, sad = util.orb_stereo_oracle(kpL[f, :cL[f]], oL[f, :cL[f]], dL[f, :cL[f]], kpR[f, :cR[f]], oR[f, :cR[f]],
                                             dR[f, :cR[f]], hL[f], hR[f], cam.baseline, cam.fx * cam.baseline)
        g_ur = out["u_right"][f, :cL[f]].cpu().numpy()
        assert np.array_equal(g_ur, ur), f
        assert np.array_equal(out["depth"][f, :cL[f]].cpu().numpy(), dp)
        assert np.array_equal(out["sad"][f, :cL[f]].cpu().numpy(), sad)
        n_match += int((ur >= 0).sum())
    assert n_match > 300
    # the per-keypoint disparity feeds A6 (stereo_slam.py:264-289): unmatched keypoints come out as 0 -> 0.1 -> dropped
    P3, valid = fsb.backproject_stereo(t(kpL[0, :cL[0]]), out["disparity"][0, :cL[0]].contiguous())
    v = valid.cpu().numpy().astype(bool)
    m = out["u_right"][0, :cL[0]].cpu().numpy() >= 0
    assert np.array_equal(v, m & v) and v.sum() > 50
