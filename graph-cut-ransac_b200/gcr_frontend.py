"""Feature front end: the reference's ``examples/utils.py`` helpers (scale_features_from_sift :5-26,
orientation_features_from_sift :29-49) on the B200 engine's C ABI (``gcr_set_keypoints``).

Same names, same arguments (a list of ``cv2.KeyPoint``-like objects with ``pt``, ``size``, ``angle``), same return
values; in addition every function accepts the keypoints as an ``[N,4]`` array of rows (x, y, size, angle in degrees) in
host memory (numpy) or device memory (a CUDA ``torch.Tensor``, read in place), which is how a batched SIFT front end
on the GPU hands them over.  ``keypoints_to_engine`` leaves the two feature arrays resident in an engine, ready for
``gcr_run`` without a host round trip.  There is no CPU path: without libgcr_b200.so / a CUDA device the calls fail.
"""
from __future__ import annotations

import numpy as np

import gcr_capi


def _as_rows(keypoints):
    """list of KeyPoint-like objects | [N,4] numpy array | CUDA tensor  ->  (host array or None, tensor or None)"""
    if hasattr(keypoints, "is_cuda"):  # torch tensor
        t = keypoints
        if t.dim() != 2 or t.shape[1] != 4:
            raise ValueError("keypoints must be an [N,4] tensor of (x, y, size, angle in degrees)")
        if not t.is_cuda:
            return t.numpy(), None
        import torch
        if t.dtype not in (torch.float32, torch.float64):
            t = t.to(torch.float64)
        return None, t.contiguous()
    if isinstance(keypoints, np.ndarray):
        return keypoints, None
    # cv2.KeyPoint stores float32 fields; Python reads them as exact doubles, and so does the float32 path of the kernel
    rows = np.array([[kp.pt[0], kp.pt[1], kp.size, kp.angle] for kp in keypoints], dtype=np.float64).reshape(-1, 4)
    return rows, None


def keypoints_to_engine(engine: gcr_capi.Engine, keypoints, want_sizes: bool = True):
    """Build both feature arrays on the device and make them the engine's resident features.
    Returns (n_scale, n_orient, sizes or None)."""
    host, dev = _as_rows(keypoints)
    if dev is not None:
        import torch
        return engine.set_keypoints(device_ptr=dev.data_ptr(), n=dev.shape[0], is_f32=dev.dtype == torch.float32,
                                    want_sizes=want_sizes)
    return engine.set_keypoints(host, want_sizes=want_sizes)


def scale_features_from_sift(keypoints, device: int = 0):
    """One (x, y, size) row per unique (int(x), int(y)): the first keypoint of each coordinate, in order of first appearance."""
    host, dev = _as_rows(keypoints)
    eng = gcr_capi.Engine(dev.device.index if dev is not None else device)
    try:
        if _n_rows(host, dev) == 0:
            return np.zeros((0, 3))
        _set(eng, host, dev, False)
        return eng.get_features()[0]
    finally:
        eng.close()


def orientation_features_from_sift(keypoints, device: int = 0):
    """(x, y, deg2rad(angle)) for every keypoint with angle != -1, and 0.5 * size per row."""
    host, dev = _as_rows(keypoints)
    eng = gcr_capi.Engine(dev.device.index if dev is not None else device)
    try:
        if _n_rows(host, dev) == 0:
            return np.zeros((0, 3)), np.zeros(0)
        _, _, sizes = _set(eng, host, dev, True)
        return eng.get_features()[1], sizes
    finally:
        eng.close()


def _n_rows(host, dev):
    return len(host) if host is not None else dev.shape[0]


def _set(eng, host, dev, want_sizes):
    try:
        return keypoints_to_engine(eng, dev if dev is not None else host, want_sizes)
    except gcr_capi.GcrError as e:
        if e.code == gcr_capi.ERR_INVALID:
            raise ValueError(str(e)) from None
        raise
