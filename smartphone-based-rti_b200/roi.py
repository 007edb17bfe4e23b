"""ROI extraction upstream of the fit, on the GPU.

Mirrors ``ut.get_ROI(static_img, static_shape_points, hsv=True)``
(Utils/utilities.py:302-342) and the V-channel pick of FeatureMatcher.py:183-184:
BGR frames in, HSV ROI / intensity frames out.  The window arithmetic (centre of the
marker corners, clamping quirks) is host code; the crop + 8-bit BGR->HSV conversion is
``rti_roi_extract``.
"""
from __future__ import annotations

import numpy as np

from . import _capi
from . import constants as _cst
from .analysis import _stream_ptr, _torch


def get_corners_center(corners, height, width):
    """Utils/utilities.py:272-299 (Python round: half to even)."""
    max_x, min_x, max_y, min_y = 0, width, 0, height
    for corner in corners:
        x, y = corner
        max_x, min_x = max(max_x, x), min(min_x, x)
        max_y, min_y = max(max_y, y), min(min_y, y)
    return round(max_x - (max_x - min_x) / 2), round(max_y - (max_y - min_y) / 2)


def roi_window(shape, static_shape_points, roi_diameter=None):
    """(x_min, x_max, y_min, y_max) of get_ROI (Utils/utilities.py:311-327); the reference's
    clamping is kept as is: a negative lower bound becomes 1, an upper bound past the edge
    becomes size - 1.  ROI_DIAMETER is read at call time like the reference does."""
    if roi_diameter is None:
        roi_diameter = _cst.ROI_DIAMETER
    radius = round(roi_diameter / 2)
    h, w = int(shape[0]), int(shape[1])
    xc, yc = get_corners_center(np.array(static_shape_points), h, w)
    x_min, x_max, y_min, y_max = xc - radius, xc + radius, yc - radius, yc + radius
    if x_min < 0:
        x_min = 1
    if x_max > w:
        x_max = w - 1
    if y_min < 0:
        y_min = 1
    if y_max > h:
        y_max = h - 1
    return int(x_min), int(x_max), int(y_min), int(y_max)


def extract_device(d_bgr, window, hsv=True, out=None):
    """d_bgr: uint8 CUDA tensor (F, H, W, 3) or (H, W, 3).  Returns a uint8 CUDA tensor
    (F, rows, cols, 3) for hsv=True or (F, rows, cols) (V only) for hsv=False; a 3-D input
    gives the result without the frame axis."""
    torch = _torch()
    single = d_bgr.dim() == 3
    if single:
        d_bgr = d_bgr[None]
    assert d_bgr.dtype == torch.uint8 and d_bgr.is_cuda and d_bgr.shape[-1] == 3 and d_bgr.is_contiguous()
    f, h, w, _ = d_bgr.shape
    x_min, x_max, y_min, y_max = window
    rows, cols = y_max - y_min, x_max - x_min
    if rows <= 0 or cols <= 0:
        raise Exception("ROI is empty")
    shape = (f, rows, cols, 3) if hsv else (f, rows, cols)
    if out is None:
        out = torch.empty(shape, dtype=torch.uint8, device=d_bgr.device)
    assert tuple(out.shape) == shape and out.is_contiguous()
    per_frame = rows * cols * (3 if hsv else 1)
    for f0 in range(0, f, 65535):
        cnt = min(65535, f - f0)
        _capi.call("rti_roi_extract", d_bgr[f0].data_ptr(), h * w * 3, cnt, h, w, x_min, y_min, rows, cols,
                   1 if hsv else 0, out[f0].data_ptr(), per_frame, _stream_ptr(torch))
    return out[0] if single else out


def get_ROI(static_img, static_shape_points, grayscale=False, hsv=False, show_roi=False):
    """Drop-in for ut.get_ROI with hsv=True (Utils/utilities.py:302-342): ndarray in,
    ndarray out.  The grayscale / plain-crop / show_roi variants are not on the path."""
    if grayscale or not hsv or show_roi:
        raise NotImplementedError("only get_ROI(..., hsv=True) is on the fitting path")
    torch = _torch()
    img = np.ascontiguousarray(static_img, dtype=np.uint8)
    window = roi_window(img.shape, static_shape_points)
    return extract_device(torch.from_numpy(img).cuda(), window, hsv=True).cpu().numpy()


def frames_to_stack(d_bgr_frames, static_shape_points):
    """Batched FeatureMatcher.py:183-184: the V-channel ROI of every BGR frame, as the
    uint8 (N, rows, cols) device stack the fit kernels read (analysis.FrameStack)."""
    window = roi_window(d_bgr_frames.shape[1:3], static_shape_points)
    return extract_device(d_bgr_frames, window, hsv=False)
