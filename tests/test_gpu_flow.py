"""GPU parity of the sparse optical flow (reference flow.py: cv2.goodFeaturesToTrack + cv2.calcOpticalFlowPyrLK with the
parameters of flow.py:11-20) against the cv2 goldens (tests/golden/flow.npz) and, when cv2 is importable on the box, against
cv2 itself on further seeds.  Corner lists are index-exact (coordinates and order); tracked points within 0.05 px."""
import numpy as np
import pytest
import torch

import synth

pytestmark = pytest.mark.gpu


def test_bgr_to_gray_exact(golden):
    from maavss_b200 import flow
    rgb = np.random.RandomState(5).randint(0, 256, (32, 40, 3), dtype=np.uint8)
    out = flow.bgr_to_gray(torch.from_numpy(rgb).cuda()).cpu().numpy()
    assert np.array_equal(out, golden("flow")["gray"])


@pytest.mark.parametrize("seed", [1, 2, 3])
def test_good_features_golden(golden, seed):
    from maavss_b200 import flow
    g = golden("flow")
    a, _ = synth.flow_frames(seed)
    ga = torch.from_numpy(a).cuda()
    p0 = flow.good_features_to_track(ga, mask=None, **flow.FEATURE_PARAMS)
    assert p0.shape == g[f"s{seed}_p0"].shape and np.array_equal(p0.cpu().numpy(), g[f"s{seed}_p0"])
    dense = flow.good_features_to_track(ga, maxCorners=100, qualityLevel=0.01, minDistance=7, blockSize=7)
    assert np.array_equal(dense.cpu().numpy(), g[f"s{seed}_dense"])


@pytest.mark.parametrize("seed", [1, 2, 3])
def test_pyr_lk_golden(golden, seed):
    from maavss_b200 import flow
    g = golden("flow")
    a, b = synth.flow_frames(seed)
    ga, gb = torch.from_numpy(a).cuda(), torch.from_numpy(b).cuda()
    p0 = torch.from_numpy(g[f"s{seed}_p0"]).cuda()
    p1, st, err = flow.calc_optical_flow_pyr_lk(ga, gb, p0, None, **flow.LK_PARAMS)
    assert np.array_equal(st.cpu().numpy(), g[f"s{seed}_st"])
    keep = g[f"s{seed}_st"].ravel() == 1
    d = np.abs(p1.cpu().numpy() - g[f"s{seed}_p1"]).reshape(-1, 2)[keep]
    print("LK max deviation from cv2 (px):", float(d.max()) if d.size else 0.0)
    assert d.size == 0 or float(d.max()) <= 0.05
    e = np.abs(err.cpu().numpy() - g[f"s{seed}_err"]).ravel()[keep]
    assert e.size == 0 or float(e.max()) <= 0.05
    # the loop body of flow.py:37-41: tracked points follow the synthetic translation (1.7, -2.3)
    old, new = flow.track(ga, gb)
    if len(old):
        mv = (new - old).cpu().numpy()
        assert np.abs(np.median(mv, 0) - np.array([1.7, -2.3])).max() < 0.3


@pytest.mark.parametrize("seed,shape,shift", [(11, (256, 256), (0.4, 0.9)), (12, (240, 320), (-3.2, 2.6)), (13, (97, 131), (5.5, -4.5)),
                                              (14, (256, 256), (12.0, 9.0))])
def test_flow_against_cv2_live(seed, shape, shift):
    """More geometries (odd sizes, shifts larger than the window at level 0, points that leave the frame) directly against cv2."""
    cv2 = pytest.importorskip("cv2")
    from maavss_b200 import flow
    a, b = synth.flow_frames(seed, H=shape[0], W=shape[1], shift=shift)
    ga, gb = torch.from_numpy(a).cuda(), torch.from_numpy(b).cuda()
    for q in (0.3, 0.02):
        ref0 = cv2.goodFeaturesToTrack(a, mask=None, maxCorners=100, qualityLevel=q, minDistance=7, blockSize=7)
        p0 = flow.good_features_to_track(ga, maxCorners=100, qualityLevel=q, minDistance=7, blockSize=7)
        assert np.array_equal(p0.cpu().numpy(), ref0), q
    # track the dense set plus points near / across the border
    extra = np.array([[[1.0, 1.0]], [[shape[1] - 2.0, shape[0] - 2.0]], [[shape[1] / 2.0, 0.5]], [[3.3, shape[0] - 1.2]]], np.float32)
    pts = np.concatenate([ref0, extra], 0)
    r1, rst, rerr = cv2.calcOpticalFlowPyrLK(a, b, pts, None, winSize=(15, 15), maxLevel=2, criteria=(3, 10, 0.03))
    p1, st, err = flow.calc_optical_flow_pyr_lk(ga, gb, torch.from_numpy(pts).cuda(), None, **flow.LK_PARAMS)
    st, p1 = st.cpu().numpy(), p1.cpu().numpy()
    agree = st.ravel() == rst.ravel()
    assert agree.mean() >= 0.97, (int((~agree).sum()), len(agree))          # a status may flip on a window at the eigenvalue threshold
    both = agree & (rst.ravel() == 1)
    d = np.abs(p1 - r1).reshape(-1, 2)[both]
    print("live: points", len(pts), "tracked", int(both.sum()), "max deviation px", float(d.max()) if d.size else 0.0)
    assert d.size == 0 or float(np.quantile(d.max(1), 0.98)) <= 0.05
