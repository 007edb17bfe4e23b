"""Headless relighting (interactive_relighting.py:11-39 without the HighGUI window):
cursor position -> light direction -> image, written to a device / host buffer."""
from __future__ import annotations

import numpy as np

from . import _capi
from .analysis import InterpolationResult, _stream_ptr, _torch


def cursor_to_light(x, y, shape):
    """draw_light_roi_position(..., to_light_vector=True) (Utils/utilities.py:368-376)."""
    h, w = shape
    lx = round(2 * (x / w) - 1, 2)
    ly = round(2 * (y / h) - 1, 2)
    if lx >= 0.99:
        lx = 0.98
    if ly >= 0.99:
        ly = 0.98
    return lx, ly


def light_to_index(l):
    """round((1 + l) / 2 * 100) (interactive_relighting.py:25-26), Python round-half-even."""
    return round((1 + l) / 2 * 100)


class Relighter:
    """Re-renders a PTM coefficient map for arbitrary light directions on the GPU.

    coef: float64 (6, P) device tensor or ndarray (an InterpolationResult also works);
    hsv_roi: optional uint8 (rows, cols, 3) HSV image whose V channel is replaced and which is
    converted HSV->BGR like cv2.cvtColor (interactive_relighting.py:37-38), including its per-row
    scalar tail (the last cols % 32 pixels of a row round where the vector body truncates).
    precision 'f64' reproduces the table values bit for bit at grid nodes; 'f32' halves the
    bytes read per frame.
    """

    def __init__(self, coef, rows, cols, hsv_roi=None, precision="f64"):
        torch = _torch()
        if isinstance(coef, InterpolationResult):
            coef = coef.coef
        if isinstance(coef, np.ndarray):
            coef = torch.from_numpy(np.ascontiguousarray(coef, dtype=np.float64).reshape(6, -1)).cuda()
        self.rows, self.cols, self.p = rows, cols, rows * cols
        assert coef.shape == (6, self.p) and coef.dtype == torch.float64
        self.precision = precision
        if precision == "f32":
            c32 = torch.empty((6, self.p), dtype=torch.float32, device="cuda")
            _capi.call("rti_coef_to_f32", coef.data_ptr(), 6 * self.p, c32.data_ptr(), _stream_ptr(torch))
            self.coef = c32
        elif precision == "f64":
            self.coef = coef.contiguous()
        else:
            raise ValueError("precision must be 'f64' or 'f32'")
        self.hsv = None
        if hsv_roi is not None:
            hsv = np.ascontiguousarray(hsv_roi, dtype=np.uint8)
            assert hsv.shape == (rows, cols, 3)
            self.hsv = torch.from_numpy(hsv).cuda()
        shape = (rows, cols, 3) if self.hsv is not None else (rows, cols)
        self.out = torch.empty(shape, dtype=torch.uint8, device="cuda")

    def render(self, lu, lv):
        """One frame for light direction (lu, lv) into self.out (device); returns it."""
        torch = _torch()
        hsv_ptr = self.hsv.data_ptr() if self.hsv is not None else None
        name = "rti_relight_ptm" if self.precision == "f64" else "rti_relight_ptm_f32"
        _capi.call(name, self.coef.data_ptr(), self.p, self.p, self.cols, float(lu), float(lv), hsv_ptr,
                   self.out.data_ptr(), _stream_ptr(torch))
        return self.out

    def render_cursor(self, x, y, shape=None):
        """relighting_event for mouse position (x, y): the light direction is snapped to the
        table grid exactly like the reference's index arithmetic."""
        shape = (self.rows, self.cols) if shape is None else shape
        lx, ly = cursor_to_light(x, y, shape)
        axis_lu = -1.0 + 0.02 * light_to_index(lx)
        axis_lv = -1.0 + 0.02 * light_to_index(ly)
        return self.render(round(axis_lu, 2), round(axis_lv, 2))


def relight_from_table(image, hsv_roi=None):
    """Table-lookup variant: clamp an int32 (rows, cols) table image to [0, 255] and
    (optionally) convert (interactive_relighting.py:33-38).  Returns a uint8 ndarray."""
    torch = _torch()
    img = torch.from_numpy(np.ascontiguousarray(image, dtype=np.int32)).cuda()
    rows, cols = img.shape
    hsv = None
    if hsv_roi is not None:
        hsv = torch.from_numpy(np.ascontiguousarray(hsv_roi, dtype=np.uint8)).cuda()
    out = torch.empty((rows, cols, 3) if hsv is not None else (rows, cols), dtype=torch.uint8, device="cuda")
    _capi.call("rti_relight_from_image", img.data_ptr(), rows * cols, cols,
               hsv.data_ptr() if hsv is not None else None, out.data_ptr(), _stream_ptr(torch))
    return out.cpu().numpy()
