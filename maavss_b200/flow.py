"""Device versions of the two OpenCV calls of the reference's flow.py (flow.py:11-20,26-37): Shi-Tomasi corner selection and
pyramidal Lucas-Kanade tracking between gray frames, through the C ABI (csrc/flow.cu).  Same argument names and return
layouts as cv2 (`(n, 1, 2)` float32 points, `(n, 1)` uint8 status, `(n, 1)` float32 error) so the body of flow.py's loop
runs unchanged on CUDA tensors; drawing and `cv2.imshow` are a GUI and left out (no graphics stack on this box)."""
import ctypes

import torch

from ._lib import check, dptr, fptr, lib, stream_ptr

# flow.py:11-20
FEATURE_PARAMS = dict(maxCorners=100, qualityLevel=0.3, minDistance=7, blockSize=7)
LK_PARAMS = dict(winSize=(15, 15), maxLevel=2, criteria=(3, 10, 0.03))      # cv2.TERM_CRITERIA_EPS | cv2.TERM_CRITERIA_COUNT = 3


def _gray(t):
    if not t.is_cuda or t.dtype != torch.uint8 or t.dim() != 2:
        raise RuntimeError("flow ops need a CUDA uint8 (H, W) gray frame (no CPU fallback)")
    return t.contiguous()


def bgr_to_gray(frame):
    """cv2.cvtColor(frame, cv2.COLOR_BGR2GRAY) for a (H, W, 3) uint8 CUDA frame."""
    f = frame.contiguous()
    if not f.is_cuda or f.dtype != torch.uint8 or f.shape[-1] != 3:
        raise RuntimeError("bgr_to_gray needs a CUDA uint8 (H, W, 3) frame")
    out = torch.empty(f.shape[:-1], device=f.device, dtype=torch.uint8)
    check(lib().maavss_bgr_to_gray(dptr(f), dptr(out), out.numel(), stream_ptr()))
    return out


def good_features_to_track(gray, mask=None, maxCorners=100, qualityLevel=0.3, minDistance=7, blockSize=7):
    """cv2.goodFeaturesToTrack -> (n, 1, 2) float32 (x, y), strongest corner first.  Reads the corner count back (one int)."""
    if mask is not None:
        raise NotImplementedError("mask is not used by flow.py")
    g = _gray(gray)
    H, W = g.shape
    L = lib()
    corners = torch.empty(maxCorners, 2, device=g.device, dtype=torch.float32)
    count = torch.zeros(1, device=g.device, dtype=torch.int32)
    nbytes = L.maavss_flow_features_workspace_bytes(H, W)
    ws = torch.empty(nbytes, device=g.device, dtype=torch.uint8)
    check(L.maavss_flow_good_features(dptr(g), H, W, int(maxCorners), float(qualityLevel), float(minDistance), int(blockSize), fptr(corners),
                                      ctypes.cast(ctypes.c_void_p(count.data_ptr()), ctypes.POINTER(ctypes.c_int)), dptr(ws), ctypes.c_longlong(nbytes),
                                      stream_ptr()))
    n = int(count.item())
    return corners[:n].reshape(n, 1, 2)


def calc_optical_flow_pyr_lk(prev_gray, next_gray, p0, p1=None, winSize=(15, 15), maxLevel=2, criteria=(3, 10, 0.03), minEigThreshold=1e-4):
    """cv2.calcOpticalFlowPyrLK -> (p1 (n, 1, 2) float32, st (n, 1) uint8, err (n, 1) float32), all on the device."""
    if p1 is not None:
        raise NotImplementedError("an initial flow estimate is not used by flow.py")
    a, b = _gray(prev_gray), _gray(next_gray)
    if a.shape != b.shape or winSize[0] != winSize[1]:
        raise RuntimeError("frames must have one shape and the window must be square")
    H, W = a.shape
    pts = p0.reshape(-1, 2).to(torch.float32).contiguous()
    n = pts.shape[0]
    nxt = torch.empty_like(pts)
    st = torch.empty(n, device=a.device, dtype=torch.uint8)
    err = torch.empty(n, device=a.device, dtype=torch.float32)
    if n == 0:
        return nxt.reshape(0, 1, 2), st.reshape(0, 1), err.reshape(0, 1)
    _, max_iter, eps = criteria
    L = lib()
    nbytes = L.maavss_flow_lk_workspace_bytes(H, W, int(maxLevel))
    ws = torch.empty(nbytes, device=a.device, dtype=torch.uint8)
    check(L.maavss_flow_pyr_lk(dptr(a), dptr(b), H, W, fptr(pts), n, None, fptr(nxt), dptr(st), fptr(err), int(winSize[0]), int(maxLevel),
                               int(min(max(max_iter, 0), 100)), float(min(max(eps, 0.0), 10.0)), float(minEigThreshold), dptr(ws),
                               ctypes.c_longlong(nbytes), stream_ptr()))
    return nxt.reshape(n, 1, 2), st.reshape(n, 1), err.reshape(n, 1)


def track(prev_gray, next_gray, p0=None):
    """One iteration of flow.py's loop (flow.py:28,37-41) with its parameters: corners of `prev_gray` (or the given points),
    their positions in `next_gray`, and the `st == 1` selection.  -> (good_old (m, 2), good_new (m, 2))."""
    if p0 is None:
        p0 = good_features_to_track(prev_gray, mask=None, **FEATURE_PARAMS)
    p1, st, _ = calc_optical_flow_pyr_lk(prev_gray, next_gray, p0, None, **LK_PARAMS)
    keep = st.reshape(-1) == 1
    return p0.reshape(-1, 2)[keep], p1.reshape(-1, 2)[keep]
