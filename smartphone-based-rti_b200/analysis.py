"""Host-side mirror of the reference's analysis.py call surface for the per-pixel
reflectance fitting path, running on librti_b200.so (hand-written sm_100a CUDA).

Same names, arguments, defaults, error messages and return shapes as the
reference (citations are into the upstream repo):

    compute_intensities(data, first_only=False)                      analysis.py:196-246
    interpolate_intensities(data, interpolate_PTM=False, first_only=False)   :321-372
    prepare_images_data(data, first_only=False)                      :375-411
    compute(video_name, from_storage, storage_filepath, interpolate_PTM,
            notification_email, debug)                               :414-482

``install(analysis_module)`` rebinds the first three inside the reference's own
module so that its unmodified ``compute()`` runs on the GPU.

Return objects index like the reference's but stay on the device until touched:
the reference's dense returns are 12 B per pixel.light and 80 kB per pixel, which
do not scale.  PyTorch is used only for device memory, streams
and (in sharding.py) torch.distributed; all arithmetic is in the CUDA library.
There is no CPU fallback.
"""
from __future__ import annotations

import os
import sys

import numpy as np

from . import _capi, table_io
from . import constants as _default_constants

_cst = _default_constants           # rebound by install() to the reference's constants module

# Numerical path of the linear-RBF mode (per-pixel lights).  "fast" (default): tile-preconditioned
# conjugate-gradient fit (rti_rbf_fit_pcg; pixels it cannot converge are refitted by the direct kernel)
# and float32-distance grid evaluation (rti_rbf_table_fast): grid values within ~2e-4 of the grid
# maximum of scipy's (north_star allows 1e-3), ~4x faster.  "fp64": direct fp64 Cholesky + fp64
# evaluation, 1e-8.  The path taken is recorded on InterpolationResult.path.
RBF_PRECISION = os.environ.get("RTI_RBF_PRECISION", "fast")
RBF_PCG_MAX_ITERS = 32
RBF_PCG_TOL = 0.01                  # max-norm node residual, in intensity units
# Linear RBF from frame records: when the three (P, N) arrays of compute_intensities plus the (P, N) weights would
# take more than this fraction of the free device memory (20 B per pixel.light: 21 GB at 1024^2 x 1000, 335 GB at
# 4096^2 x 1000), nothing is materialised for the whole ROI: gather, fit and evaluation run per band of ROI rows
# while the table is produced (same numbers; the weights of a pixel are recomputed if res[y][x] asks for them).
RBF_BANDED_FRACTION = 0.25
# Device bytes of one table tile buffer (two are used): prepare_images_data / write_table_file evaluate
# the [ly][lx] images in ROW TILES of the ROI through these, so the (G, G, P) table never has to fit on
# the device (42 GB at 1024^2, 671 GB at 4096^2).
TABLE_TILE_BYTES = 1 << 30


def _torch():
    import torch
    if not torch.cuda.is_available():
        raise RuntimeError("smartphone-based-rti_b200 needs a CUDA device (sm_100a); no CPU fallback")
    return torch


def _stream_ptr(torch):
    return torch.cuda.current_stream().cuda_stream


def light_axis(step=None) -> np.ndarray:
    """One axis of the light grid: np.around(np.mgrid[-1:1:step], 2) (analysis.py:345-347)."""
    step = _cst.INTERPOLATION_PARAM if step is None else step
    return np.around(np.mgrid[-1:1:step], decimals=2)


# --------------------------------------------------------------------------- records
def _stack_frames(data, rows, cols):
    """list[(uint8 (H,W), float64 (3,))] -> contiguous (N, rows, cols) uint8 + (N, 3) float64."""
    n = len(data)
    frames = np.empty((n, rows, cols), dtype=np.uint8)
    cams = np.empty((n, 3), dtype=np.float64)
    for i, rec in enumerate(data):
        roi = np.asarray(rec[0])
        if roi.ndim != 2 or roi.shape[0] < rows or roi.shape[1] < cols:
            # the reference raises IndexError at intensities[y][x] (analysis.py:232)
            raise IndexError("frame %d ROI has shape %s, need at least (%d, %d)" % (i, roi.shape, rows, cols))
        from .ingest import roi_as_uint8
        frames[i] = roi_as_uint8(roi[:rows, :cols], i)
        cams[i] = np.asarray(rec[1], dtype=np.float64).reshape(3)
    return frames, cams


class FrameStack:
    """Device-resident frame records: uint8 (N, rows, cols) + float64 (N, 3), plus the
    global row of the slab's first row (row0) for row-tile sharding."""

    def __init__(self, frames, cams, row0=0):
        torch = _torch()
        frames = np.ascontiguousarray(frames, dtype=np.uint8)
        cams = np.ascontiguousarray(cams, dtype=np.float64)
        self.n, self.rows, self.cols = frames.shape
        self.row0 = int(row0)
        self.d_frames = torch.from_numpy(frames).cuda(non_blocking=False)
        self.d_cams = torch.from_numpy(cams).cuda()

    @classmethod
    def from_device(cls, d_frames, d_cams, row0=0):
        self = cls.__new__(cls)
        self.n, self.rows, self.cols = d_frames.shape
        self.row0 = int(row0)
        self.d_frames, self.d_cams = d_frames.contiguous(), d_cams.contiguous()
        return self

    @property
    def n_pixels(self):
        return self.rows * self.cols


class _SoA:
    """The three (rows, cols, N) arrays of compute_intensities on the device, produced
    on first use by the gather kernel."""

    def __init__(self, stack: FrameStack):
        self.stack = stack
        self._arrays = None

    def device_arrays(self):
        if self._arrays is None:
            torch = _torch()
            s = self.stack
            shape = (s.n_pixels, s.n)
            lx = torch.empty(shape, dtype=torch.float32, device="cuda")
            ly = torch.empty(shape, dtype=torch.float32, device="cuda")
            inten = torch.empty(shape, dtype=torch.int32, device="cuda")
            _capi.call("rti_gather_soa", s.d_frames.data_ptr(), s.rows * s.cols, s.d_cams.data_ptr(),
                       s.n, s.row0, s.rows, s.cols, lx.data_ptr(), ly.data_ptr(), inten.data_ptr(),
                       _stream_ptr(torch))
            self._arrays = (lx, ly, inten)
        return self._arrays


class DeviceArrayView:
    """One of pixels_lx / pixels_ly / pixels_intensity: indexes and converts like the
    (rows, cols, N) ndarray the reference returns (analysis.py:217-219, 245)."""

    def __init__(self, soa: _SoA, which: int):
        self._soa, self._which = soa, which
        s = soa.stack
        self.shape = (s.rows, s.cols, s.n)
        self.dtype = np.dtype(np.int32 if which == 2 else np.float32)
        self.ndim = 3

    def __len__(self):
        return self.shape[0]

    def device(self):
        return self._soa.device_arrays()[self._which].view(self.shape)

    def __array__(self, dtype=None, copy=None):
        a = self.device().cpu().numpy()
        return a if dtype is None else a.astype(dtype)

    def __getitem__(self, idx):
        out = self.device()[idx]
        return out.cpu().numpy() if out.ndim else out.item()


def compute_intensities(data, first_only=False):
    """Compute light vectors intensities foreach frame pixel (analysis.py:196-246).

    :param data: list of (intensities uint8 (R,R), camera_position float64 (3,)) per frame
    :param first_only: only pixel (0, 0)
    :return: (pixels_lx, pixels_ly, pixels_intensity), each indexable [y][x][i]
    """
    if data is None or len(data) <= 0:
        raise Exception("Error computing intensities: results are empty")
    print("Computing intensities values:")
    range_val = 1 if first_only else _cst.ROI_DIAMETER
    torch = _torch()
    from . import ingest
    frames, cams = ingest.pinned_stack(data, range_val, range_val)      # one copy, into pinned memory
    stack = FrameStack.from_device(frames.cuda(non_blocking=True), torch.from_numpy(cams).cuda())
    stack._pinned = frames                                               # alive until the upload has run
    soa = _SoA(stack)
    return (DeviceArrayView(soa, 0), DeviceArrayView(soa, 1), DeviceArrayView(soa, 2))


# --------------------------------------------------------------------- interpolation
class InterpolationResult:
    """interpolated_intensities[y][x] -> (G, G) float64 indexed [ly][lx]
    (analysis.py:349, 372), backed by the per-pixel model on the device:
    PTM: coefficient map float64 (6, P);  RBF: nodes (lx, ly) and weights float64 (P, N).
    A pixel's grid is evaluated on demand."""

    def __init__(self, mode, rows, cols, axis, coef=None, lx=None, ly=None, weights=None, row0=0, shared=None,
                 stack=None, precision="fp64", path=None):
        self.mode, self.rows, self.cols, self.row0 = mode, rows, cols, row0
        self.axis = np.ascontiguousarray(axis, dtype=np.float64)
        self._coef, self.lx, self.ly, self.weights = coef, lx, ly, weights
        self.shared = shared            # mode "rbf_shared": a SharedLightRbf
        self.stack = stack              # mode "ptm" from frame records: coefficients are fitted on demand, or by
        self._coef_done = coef is not None   # the fused fit + table kernel while the table is produced
        self.precision = precision      # "fast" | "fp64" (RBF modes)
        self.fp64_idx = None            # "fast" RBF: pixel indices (device int64) that are evaluated in fp64 anyway
        self._band = None               # banded RBF (stack-backed, no whole-ROI arrays): the last fitted band
        self.path = path or mode        # which kernels produced / will produce the numbers
        self._d_axis = None

    @property
    def coef(self):
        """PTM coefficient map float64 (6, P) on the device (fitted on first use when stack-backed)."""
        if self.mode == "ptm" and not self._coef_done:
            coef, bad = _fit_ptm_from_stack(self.stack, out=self._coef)
            _raise_if_bad(bad, "PTM")
            self._coef, self._coef_done = coef, True
        return self._coef

    @property
    def n_pixels(self):
        return self.rows * self.cols

    def d_axis(self):
        if self._d_axis is None:
            self._d_axis = _torch().from_numpy(self.axis).cuda()
        return self._d_axis

    def __len__(self):
        return self.rows

    def __getitem__(self, y):
        if not -self.rows <= y < self.rows:
            raise IndexError("list index out of range")
        return _ResultRow(self, y % self.rows)

    @property
    def banded(self):
        return self.mode == "rbf" and self.weights is None and self.stack is not None

    def rbf_band(self, r0, tr):
        """(lx, ly, weights, fp64_idx) of ROI rows [r0, r0 + tr) of a banded RBF result: gather + fit of just
        those rows (rti_gather_soa on the frame rows, then the pcg or the direct fit); the last band is kept."""
        if self._band is not None and self._band[0] == (r0, tr):
            return self._band[1]
        torch = _torch()
        st = self.stack
        tp = tr * st.cols
        shape = (tp, st.n)
        lx = torch.empty(shape, dtype=torch.float32, device="cuda")
        ly = torch.empty(shape, dtype=torch.float32, device="cuda")
        inten = torch.empty(shape, dtype=torch.int32, device="cuda")
        _capi.call("rti_gather_soa", st.d_frames.data_ptr() + r0 * st.cols, st.n_pixels, st.d_cams.data_ptr(), st.n,
                   st.row0 + r0, tr, st.cols, lx.data_ptr(), ly.data_ptr(), inten.data_ptr(), _stream_ptr(torch))
        bad = torch.zeros(1, dtype=torch.int32, device="cuda")
        if self.precision == "fast" and 3 <= st.n <= 1025:
            weights, info = rbf_fit_device_pcg(lx, ly, inten, tr, st.cols, bad)
            idx = info["fp64_idx"]
        else:
            weights, idx = rbf_fit_device(lx, ly, inten, bad), None
        _raise_if_bad(bad, "RBF")
        del inten
        self._band = ((r0, tr), (lx, ly, weights, idx))
        return self._band[1]

    def grids(self, p0, count):
        """float64 (count, G, G) grids of pixels [p0, p0 + count)."""
        torch = _torch()
        g = self.axis.size
        out = torch.empty((count, g, g), dtype=torch.float64, device="cuda")
        if self.banded:
            # the 16-row band(s) that hold the pixels (tile-aligned, so the fit is the one the table gets)
            done = 0
            while done < count:
                p = p0 + done
                y = p // self.cols
                b0 = (y // 16) * 16
                tr = min(16, self.rows - b0)
                lx, ly, w, _ = self.rbf_band(b0, tr)
                q0 = p - b0 * self.cols
                c = min(count - done, tr * self.cols - q0)
                n = w.shape[1]
                _capi.call("rti_rbf_grids", lx.data_ptr() + 4 * q0 * n, ly.data_ptr() + 4 * q0 * n,
                           w.data_ptr() + 8 * q0 * n, c, n, self.d_axis().data_ptr(), g,
                           out.data_ptr() + 8 * done * g * g, _stream_ptr(torch))
                done += c
            return out.cpu().numpy()
        if self.mode == "ptm":
            _capi.call("rti_ptm_grids", self.coef.data_ptr() + 8 * p0, self.n_pixels, count,
                       self.d_axis().data_ptr(), g, out.data_ptr(), _stream_ptr(torch))
        elif self.mode == "rbf_shared":
            out = self.shared.grids(p0, count).view(count, g, g)
        else:
            n = self.weights.shape[1]
            _capi.call("rti_rbf_grids", self.lx.data_ptr() + 4 * p0 * n, self.ly.data_ptr() + 4 * p0 * n,
                       self.weights.data_ptr() + 8 * p0 * n, count, n, self.d_axis().data_ptr(), g,
                       out.data_ptr(), _stream_ptr(torch))
        return out.cpu().numpy()

    def coefficients(self) -> np.ndarray:
        """PTM only: (6, rows, cols) float64 a_matrix per pixel (analysis.py:298)."""
        if self.mode != "ptm":
            raise ValueError("coefficients() is defined for PTM results")
        return self.coef.cpu().numpy().reshape(6, self.rows, self.cols)

    def to_list(self):
        """Materialise the reference's nested list (80 kB per pixel!)."""
        out = []
        for y in range(self.rows):
            g = self.grids(y * self.cols, self.cols)
            out.append([g[x] for x in range(self.cols)])
        return out


class _ResultRow:
    def __init__(self, res, y):
        self._res, self._y = res, y

    def __len__(self):
        return self._res.cols

    def __getitem__(self, x):
        cols = self._res.cols
        if not -cols <= x < cols:
            raise IndexError("list index out of range")
        return self._res.grids(self._y * cols + x % cols, 1)[0]


def _fit_ptm_from_stack(stack: FrameStack, out=None):
    torch = _torch()
    coef = torch.empty((6, stack.n_pixels), dtype=torch.float64, device="cuda") if out is None else out
    bad = torch.zeros(1, dtype=torch.int32, device="cuda")
    _capi.call("rti_ptm_fit_frames", stack.d_frames.data_ptr(), stack.rows * stack.cols,
               stack.d_cams.data_ptr(), stack.n, stack.row0, stack.rows, stack.cols,
               coef.data_ptr(), stack.n_pixels, bad.data_ptr(), _stream_ptr(torch))
    return coef, bad


def interpolate_ptm_masked(data, valid_range, first_only=False):
    """PTM with per-pixel validity masks (extension): like
    ``interpolate_intensities(data, interpolate_PTM=True)`` but a frame contributes to a pixel only
    when its intensity there lies in ``valid_range = (lo, hi)`` (shadow / highlight rejection).  Equals
    the reference's ``_interpolate_PTM`` called with each pixel's masked sub-arrays.  ``data``: the frame
    records, a FrameStack, or the tuple compute_intensities returned.  Returns (result, valid_count)
    with valid_count an int32 (rows, cols) device tensor; pixels with fewer than 6 valid samples raise
    RtiSingularError (the reference fails in np.dot there, analysis.py:296-298)."""
    torch = _torch()
    lo, hi = int(valid_range[0]), int(valid_range[1])
    range_val = 1 if first_only else _cst.ROI_DIAMETER
    if isinstance(data, tuple) and len(data) == 3 and isinstance(data[0], DeviceArrayView):
        stack = data[0]._soa.stack
    elif isinstance(data, FrameStack):
        stack = data
    else:
        if data is None or len(data) <= 0:
            raise Exception("Error computing interpolation: results are empty or invalid")
        stack = FrameStack(*_stack_frames(data, range_val, range_val))
    coef = torch.empty((6, stack.n_pixels), dtype=torch.float64, device="cuda")
    count = torch.empty((stack.rows, stack.cols), dtype=torch.int32, device="cuda")
    bad = torch.zeros(1, dtype=torch.int32, device="cuda")
    _capi.call("rti_ptm_fit_frames_masked", stack.d_frames.data_ptr(), stack.rows * stack.cols,
               stack.d_cams.data_ptr(), stack.n, stack.row0, stack.rows, stack.cols, lo, hi,
               coef.data_ptr(), stack.n_pixels, count.data_ptr(), bad.data_ptr(), _stream_ptr(torch))
    _raise_if_bad(bad, "masked PTM")
    return InterpolationResult("ptm", stack.rows, stack.cols, light_axis(), coef=coef, row0=stack.row0), count


def _as_device_soa(data, range_val):
    """(lx, ly, inten) as device (P, N) tensors restricted to [range_val, range_val]."""
    torch = _torch()
    if all(isinstance(a, DeviceArrayView) for a in data) and data[0].shape[:2] == (range_val, range_val):
        return data[0]._soa.device_arrays()
    out = []
    for a, dt in zip(data, (np.float32, np.float32, np.int32)):
        arr = np.asarray(a)
        if arr.ndim != 3 or arr.shape[0] < range_val or arr.shape[1] < range_val:
            raise IndexError("pixel data of shape %s does not cover a %dx%d ROI" % (arr.shape, range_val, range_val))
        arr = np.ascontiguousarray(arr[:range_val, :range_val], dtype=dt)
        out.append(torch.from_numpy(arr).cuda().view(range_val * range_val, arr.shape[2]))
    return tuple(out)


class SharedLightRbf:
    """Linear RBF when every pixel has the same n nodes: one operator for all pixels
    (rti_rbf_shared_operator) applied to the packed intensities on the tensor cores."""

    def __init__(self, lx1, ly1, axis, d_axis):
        torch = _torch()
        self.lib = _capi.load()
        self.lx1, self.ly1 = lx1.contiguous(), ly1.contiguous()          # float32 (n,) on the device
        self.n = int(lx1.numel())
        self.g = int(axis.size)
        self.d_axis = d_axis
        self._ops = {}
        self.packed, self.n_pixels = None, 0
        self.bad = torch.zeros(1, dtype=torch.int32, device="cuda")

    def operator(self, kind):
        """(hi, lo, inv_scale, rows) of the grid (kind 0) or weight (kind 1) operator."""
        if kind not in self._ops:
            torch = _torch()
            rows = self.g * self.g if kind == 0 else self.n
            elems = self.lib.rti_rbf_shared_operator_elems(rows, self.n)
            hi = torch.empty(elems, dtype=torch.int16, device="cuda")
            lo = torch.empty(elems, dtype=torch.int16, device="cuda")
            inv = torch.empty(rows, dtype=torch.float32, device="cuda")
            _capi.call("rti_rbf_shared_operator", self.lx1.data_ptr(), self.ly1.data_ptr(), self.n,
                       self.d_axis.data_ptr(), self.g, kind, hi.data_ptr(), lo.data_ptr(), inv.data_ptr(),
                       self.bad.data_ptr(), _stream_ptr(torch))
            _raise_if_bad(self.bad, "RBF")
            self._ops[kind] = (hi, lo, inv, rows)
        return self._ops[kind]

    def _new_packed(self, n_pixels):
        torch = _torch()
        return torch.empty(self.lib.rti_rbf_shared_packed_elems(self.n, n_pixels), dtype=torch.int16, device="cuda")

    def pack_frames(self, d_frames, n_pixels):
        """uint8 frames (n, n_pixels) on the device -> packed fp16 intensities."""
        self.packed, self.n_pixels = self._new_packed(n_pixels), n_pixels
        _capi.call("rti_rbf_shared_pack_frames", d_frames.data_ptr(), n_pixels, self.n, n_pixels,
                   self.packed.data_ptr(), _stream_ptr(_torch()))

    def pack_soa(self, inten):
        """int32 intensities (n_pixels, n) on the device -> packed fp16; False if not exact in fp16."""
        torch = _torch()
        n_pixels = int(inten.shape[0])
        packed = self._new_packed(n_pixels)
        flag = torch.zeros(1, dtype=torch.int32, device="cuda")
        _capi.call("rti_rbf_shared_pack_soa", inten.data_ptr(), n_pixels, self.n, packed.data_ptr(),
                   flag.data_ptr(), _stream_ptr(torch))
        if int(flag.item()) > 0:
            return False
        self.packed, self.n_pixels = packed, n_pixels
        return True

    def _block(self, p0, count):
        """Packed sub-range: whole 256-pixel blocks only (the packed layout is blocked)."""
        if p0 % 256 != 0:
            raise ValueError("shared-light sub-ranges start on a 256-pixel block")
        return self.packed.data_ptr() + 2 * p0 * self.n

    def table(self, out, image_stride, bad):
        hi, lo, inv, rows = self.operator(0)
        _capi.call("rti_rbf_shared_table", self.packed.data_ptr(), self.n, self.n_pixels, hi.data_ptr(),
                   lo.data_ptr(), inv.data_ptr(), rows, out.data_ptr(), image_stride, bad.data_ptr(),
                   _stream_ptr(_torch()))

    def _apply(self, kind, p0, count):
        torch = _torch()
        hi, lo, inv, rows = self.operator(kind)
        b0 = (p0 // 256) * 256
        cnt = min(self.n_pixels, p0 + count) - b0
        out = torch.empty((cnt, rows), dtype=torch.float64, device="cuda")
        _capi.call("rti_rbf_shared_apply_f64", self._block(b0, cnt), self.n, cnt, hi.data_ptr(), lo.data_ptr(),
                   inv.data_ptr(), rows, out.data_ptr(), _stream_ptr(torch))
        return out[p0 - b0:p0 - b0 + count]

    def grids(self, p0, count):
        """float64 (count, g*g) grid values of pixels [p0, p0 + count)."""
        return self._apply(0, p0, count)

    def weights(self, p0=0, count=None):
        """float64 (count, n) RBF weights (scipy Rbf.nodes) of pixels [p0, p0 + count)."""
        return self._apply(1, p0, self.n_pixels - p0 if count is None else count)


def lights_are_shared(lx, ly):
    """True when every pixel row of the device tensors lx, ly (P, n) equals pixel 0's."""
    torch = _torch()
    n_pixels, n = lx.shape
    if n_pixels < 2:
        return False                       # a single pixel: nothing to share
    # cheap rejection first: the per-pixel data of compute_intensities differs at once
    if not (torch.equal(lx[0], lx[n_pixels - 1]) and torch.equal(ly[0], ly[n_pixels - 1])):
        return False
    flag = torch.zeros(1, dtype=torch.int32, device="cuda")
    _capi.call("rti_lights_shared", lx.data_ptr(), ly.data_ptr(), n_pixels, n, flag.data_ptr(), _stream_ptr(torch))
    return int(flag.item()) == 0


def interpolate_shared_light(data, lx, ly, interpolate_PTM=False):
    """Shared-light entry: frame records as for compute_intensities plus ONE light vector per
    frame (lx[i], ly[i]) valid for the whole ROI (distant light).  Same result object as
    interpolate_intensities would give for pixels_lx[y][x][:] = lx, pixels_ly[y][x][:] = ly,
    without materialising the (R, R, N) arrays: PTM is coef = pinv . I (rti_ptm_fit_shared),
    linear RBF one tensor-core contraction (rti_rbf_shared_*)."""
    if data is None or len(data) <= 0:
        raise Exception("Error computing interpolation: results are empty or invalid")
    torch = _torch()
    range_val = _cst.ROI_DIAMETER
    stack = data if isinstance(data, FrameStack) else FrameStack(*_stack_frames(data, range_val, range_val))
    lx1 = torch.from_numpy(np.ascontiguousarray(lx, dtype=np.float32).reshape(-1)).cuda()
    ly1 = torch.from_numpy(np.ascontiguousarray(ly, dtype=np.float32).reshape(-1)).cuda()
    if lx1.numel() != stack.n or ly1.numel() != stack.n:
        raise ValueError("need one (lx, ly) per frame: %d frames, %d / %d values" % (stack.n, lx1.numel(), ly1.numel()))
    axis = light_axis()
    p = stack.n_pixels
    if interpolate_PTM:
        pinv = torch.empty((6, stack.n), dtype=torch.float64, device="cuda")
        bad = torch.zeros(1, dtype=torch.int32, device="cuda")
        _capi.call("rti_ptm_shared_pinv", lx1.data_ptr(), ly1.data_ptr(), stack.n, pinv.data_ptr(), bad.data_ptr(),
                   _stream_ptr(torch))
        _raise_if_bad(bad, "PTM")
        coef = torch.empty((6, p), dtype=torch.float64, device="cuda")
        _capi.call("rti_ptm_fit_shared", stack.d_frames.data_ptr(), p, pinv.data_ptr(), stack.n, p,
                   coef.data_ptr(), p, _stream_ptr(torch))
        return InterpolationResult("ptm", stack.rows, stack.cols, axis, coef=coef, row0=stack.row0)
    res = InterpolationResult("rbf_shared", stack.rows, stack.cols, axis, row0=stack.row0)
    res.shared = SharedLightRbf(lx1, ly1, axis, res.d_axis())
    res.shared.pack_frames(stack.d_frames, p)
    res.shared.operator(0)                 # raises RtiSingularError on duplicate nodes, like scipy's LinAlgError
    return res


def interpolate_intensities(data, interpolate_PTM=False, first_only=False):
    """Interpolate pixel intensities (analysis.py:321-372).

    :param data: (pixels_lx, pixels_ly, pixels_intensity) as returned by compute_intensities
                 (device views) or as plain (R, R, N) ndarrays
    :param interpolate_PTM: PTM if True, else linear RBF
    :param first_only: only pixel (0, 0)
    :return: interpolated_intensities[y][x][ly][lx] => intensity (lazy, see InterpolationResult)
    """
    if data is None or len(data) != 3:
        raise Exception("Error computing interpolation: results are empty or invalid")
    print("Computing interpolation values:")
    torch = _torch()
    range_val = 1 if first_only else _cst.ROI_DIAMETER
    axis = light_axis()

    fused = (interpolate_PTM and all(isinstance(a, DeviceArrayView) for a in data)
             and data[0]._soa is data[1]._soa is data[2]._soa
             and data[0].shape[:2] == (range_val, range_val))
    if fused:
        # straight from the uint8 frame records: nothing runs yet.  prepare_images_data / write_table_file fit
        # and evaluate row tile by row tile with the fused kernel (rti_ptm_fit_table_frames); touching
        # res[y][x] or res.coefficients() first fits with rti_ptm_fit_frames instead.
        stack = data[0]._soa.stack
        if stack.n < 6:
            raise ValueError("PTM needs >= 6 light samples, got %d (analysis.py:296-298 fails in np.dot)" % stack.n)
        return InterpolationResult("ptm", range_val, range_val, axis, row0=stack.row0, stack=stack,
                                   path="ptm: fused gather+fit+table per row tile (rti_ptm_fit_table_frames)")

    from_records = (all(isinstance(a, DeviceArrayView) for a in data) and data[0]._soa is data[1]._soa is data[2]._soa
                    and data[0].shape[:2] == (range_val, range_val))
    if not interpolate_PTM and from_records and data[0]._soa._arrays is None:
        stack = data[0]._soa.stack
        free, _total = torch.cuda.mem_get_info()
        if 20.0 * stack.n_pixels * stack.n > RBF_BANDED_FRACTION * free:
            precision = RBF_PRECISION if RBF_PRECISION in ("fast", "fp64") else "fast"
            return InterpolationResult("rbf", range_val, range_val, axis, row0=stack.row0, stack=stack,
                                       precision=precision,
                                       path="rbf: banded (gather + %s fit + table per band of ROI rows; nothing "
                                            "materialised for the whole ROI)" % ("pcg" if precision == "fast" else "direct"))
    lx, ly, inten = _as_device_soa(data, range_val)
    n_pixels, n = lx.shape
    bad = torch.zeros(1, dtype=torch.int32, device="cuda")
    if interpolate_PTM:
        coef = torch.empty((6, n_pixels), dtype=torch.float64, device="cuda")
        _capi.call("rti_ptm_fit_soa", lx.data_ptr(), ly.data_ptr(), inten.data_ptr(), n_pixels, n,
                   coef.data_ptr(), n_pixels, bad.data_ptr(), _stream_ptr(torch))
        _raise_if_bad(bad, "PTM")
        return InterpolationResult("ptm", range_val, range_val, axis, coef=coef)
    precision = RBF_PRECISION if RBF_PRECISION in ("fast", "fp64") else "fast"
    if precision == "fast" and lights_are_shared(lx, ly):
        # one light vector per frame for the whole ROI: one operator, one tensor-core contraction (fp16 hi+lo
        # operator: grid values within ~2e-5 of the maximum; RBF_PRECISION = "fp64" keeps such data on the
        # per-pixel fp64 kernels)
        res = InterpolationResult("rbf_shared", range_val, range_val, axis, precision="fast",
                                  path="rbf: shared lights, one tcgen05 contraction (fp16 hi+lo operator)")
        res.shared = SharedLightRbf(lx[0].clone(), ly[0].clone(), axis, res.d_axis())
        if res.shared.pack_soa(inten):     # intensities exact in fp16 (always true for uint8 frames)
            res.shared.operator(0)
            return res
    if precision == "fast" and 3 <= n <= 1025:
        weights, info = rbf_fit_device_pcg(lx, ly, inten, range_val, range_val, bad)
        fp64_idx = info["fp64_idx"]
        path = ("rbf: pcg fit + float32-distance table; %d of %d pixels (not converged, or weights too large for "
                "float32 distances) fitted and evaluated in fp64" % (info["refit"], n_pixels))
    else:
        weights = rbf_fit_device(lx, ly, inten, bad)
        fp64_idx = None
        precision = "fp64"                 # (n outside the pcg kernel's range: everything on the fp64 kernels)
        path = "rbf: direct fp64 fit + fp64 table"
    _raise_if_bad(bad, "RBF")
    res = InterpolationResult("rbf", range_val, range_val, axis, lx=lx, ly=ly, weights=weights,
                              precision=precision, path=path)
    res.fp64_idx = fp64_idx
    return res


def rbf_fit_device(lx, ly, inten, bad, out=None):
    """Linear-RBF weights (P, N) float64 for device SoA tensors (scipy Rbf.__init__), direct fp64 solve."""
    torch = _torch()
    n_pixels, n = lx.shape
    weights = torch.empty((n_pixels, n), dtype=torch.float64, device="cuda") if out is None else out
    free, _total = torch.cuda.mem_get_info()
    lib = _capi.load()
    per_pixel = max(1, lib.rti_rbf_workspace_bytes(n, 1))
    batch = int(max(1, min(n_pixels, (free // 2) // per_pixel, 148 * 64)))
    work_bytes = lib.rti_rbf_workspace_bytes(n, batch)
    work = torch.empty(max(work_bytes, 8), dtype=torch.uint8, device="cuda")
    _capi.call("rti_rbf_fit_soa", lx.data_ptr(), ly.data_ptr(), inten.data_ptr(), n_pixels, n,
               weights.data_ptr(), work.data_ptr(), work_bytes, batch, bad.data_ptr(), _stream_ptr(torch))
    return weights


def rbf_fit_device_pcg(lx, ly, inten, rows, cols, bad, max_iters=None, tol=None):
    """Linear-RBF weights of a rows x cols slab by the tile-preconditioned conjugate gradients
    (rti_rbf_fit_pcg); pixels that did not converge (neighbouring pixels' nodes too different from the tile
    centre's: data that does not come from compute_intensities) and the pixels of tiles whose centre is
    singular are refitted with the direct fp64 kernel, which also reports genuinely singular pixels in `bad`.
    Returns (weights, info)."""
    torch = _torch()
    lib = _capi.load()
    n_pixels, n = lx.shape
    assert n_pixels == rows * cols
    max_iters = RBF_PCG_MAX_ITERS if max_iters is None else max_iters
    tol = RBF_PCG_TOL if tol is None else tol
    weights = torch.empty((n_pixels, n), dtype=torch.float64, device="cuda")
    tiles = ((rows + 15) // 16) * ((cols + 15) // 16)
    free, _total = torch.cuda.mem_get_info()
    per_tile = lib.rti_rbf_pcg_workspace_bytes(n, 2) - lib.rti_rbf_pcg_workspace_bytes(n, 1)
    in_flight = int(max(1, min(tiles, (free // 2) // max(per_tile, 1), 4096)))
    work_bytes = lib.rti_rbf_pcg_workspace_bytes(n, in_flight)
    work = torch.empty(work_bytes, dtype=torch.uint8, device="cuda")
    iters = torch.empty(n_pixels, dtype=torch.int32, device="cuda")
    pbad = torch.zeros(2, dtype=torch.int32, device="cuda")
    _capi.call("rti_rbf_fit_pcg", lx.data_ptr(), ly.data_ptr(), inten.data_ptr(), rows, cols, n, weights.data_ptr(),
               work.data_ptr(), work_bytes, max_iters, float(tol), iters.data_ptr(), pbad.data_ptr(),
               _stream_ptr(torch))
    del work
    singular_tiles, unconverged = (int(v) for v in pbad.tolist())
    info = {"singular_tiles": singular_tiles, "refit": 0, "max_iters": max_iters, "tol": tol, "fp64_idx": None}
    if singular_tiles or unconverged:
        redo = iters <= 0 if singular_tiles == 0 else (iters <= 0) | ~torch.isfinite(weights).all(dim=1)
        if singular_tiles:
            # a singular tile centre leaves its whole tile without a preconditioner
            redo |= iters >= max_iters
        idx = redo.nonzero().view(-1)
        info["refit"] = int(idx.numel())
        info["fp64_idx"] = idx             # these pixels are also EVALUATED in fp64 (table_tiles)
        if idx.numel():
            w_sub = rbf_fit_device(lx[idx].contiguous(), ly[idx].contiguous(), inten[idx].contiguous(), bad)
            weights[idx] = w_sub
    info["steps_mean"] = float(iters.abs().float().mean().item())
    return weights, info


def _raise_if_bad(bad, what):
    count = int(bad.item())          # synchronises: numerical failures surface here
    if count > 0:
        raise _capi.RtiSingularError(_capi.RTI_ERR_SINGULAR,
                                     "%s: %d pixel(s) have a singular system" % (what, count))


# ----------------------------------------------------------------------------- table
def _table_rows(res: InterpolationResult, r0, tr, out, bad):
    """Rows [r0, r0 + tr) of every [ly][lx] image into out, an int32 (G*G, tr*cols) device tensor
    (bad: int32[2] device counters: [0] singular PTM pixels, [1] values that do not fit int32)."""
    torch = _torch()
    g = res.axis.size
    cols = res.cols
    p0, tp = r0 * cols, tr * cols
    S = _stream_ptr(torch)
    if res.mode == "ptm" and not res._coef_done and res.stack is not None:
        st = res.stack
        if res._coef is None:
            res._coef = torch.empty((6, st.n_pixels), dtype=torch.float64, device="cuda")
        # fused gather + fit + table of this row tile; the coefficient map fills in as a by-product
        _capi.call("rti_ptm_fit_table_frames", st.d_frames.data_ptr() + p0, st.n_pixels, st.d_cams.data_ptr(), st.n,
                   st.row0 + r0, tr, cols, res.d_axis().data_ptr(), g, res._coef.data_ptr() + 8 * p0, st.n_pixels,
                   out.data_ptr(), tp, bad.data_ptr(), S)
        if r0 + tr >= res.rows:
            res._coef_done = True          # (a caller that stops early refits on demand)
    elif res.mode == "ptm":
        _capi.call("rti_ptm_table", res.coef.data_ptr() + 8 * p0, res.n_pixels, tp, res.d_axis().data_ptr(), g,
                   out.data_ptr(), tp, bad.data_ptr() + 4, S)
    elif res.mode == "rbf_shared":
        if (r0, tr) != (0, res.rows):
            raise ValueError("shared-light tables are produced in one piece")
        res.shared.table(out, tp, bad[1:])
    elif res.banded:
        lx, ly, w, idx = res.rbf_band(r0, tr)
        rbf_table_device(lx, ly, w, res.d_axis(), g, out, bad[1:], precision=res.precision, fp64_idx=idx)
    else:
        rbf_table_device(res.lx, res.ly, res.weights, res.d_axis(), g, out, bad[1:], p0=p0, count=tp,
                         precision=res.precision, fp64_idx=res.fp64_idx)


def rbf_table_device(lx, ly, weights, d_axis, g, out, bad, p0=0, count=None, precision="fast", fp64_idx=None):
    """int32 table columns of pixels [p0, p0 + count) of a per-pixel linear-RBF model into out (g*g, count):
    float32-distance kernel for precision "fast", except the pixels listed in fp64_idx (weights too large for
    float32 distances, see rbf_fit_device_pcg), which are evaluated by the fp64 kernel and scattered in."""
    torch = _torch()
    S = _stream_ptr(torch)
    n = weights.shape[1]
    count = lx.shape[0] - p0 if count is None else count
    fn = "rti_rbf_table_fast" if precision == "fast" else "rti_rbf_table"
    _capi.call(fn, lx.data_ptr() + 4 * p0 * n, ly.data_ptr() + 4 * p0 * n, weights.data_ptr() + 8 * p0 * n, count, n,
               d_axis.data_ptr(), g, out.data_ptr(), count, bad.data_ptr(), S)
    if precision == "fast" and fp64_idx is not None and fp64_idx.numel():
        sel = fp64_idx[(fp64_idx >= p0) & (fp64_idx < p0 + count)]
        k = int(sel.numel())
        if k:
            sub = torch.empty((g * g, k), dtype=torch.int32, device="cuda")
            lx_s, ly_s, w_s = lx[sel].contiguous(), ly[sel].contiguous(), weights[sel].contiguous()
            _capi.call("rti_rbf_table", lx_s.data_ptr(), ly_s.data_ptr(), w_s.data_ptr(), k, n, d_axis.data_ptr(), g,
                       sub.data_ptr(), k, bad.data_ptr(), S)
            out[:, sel - p0] = sub


def _raise_if_bad_table(bad):
    singular, unrepresentable = (int(v) for v in bad.tolist())       # synchronises
    if singular > 0:
        raise _capi.RtiSingularError(_capi.RTI_ERR_SINGULAR, "PTM: %d pixel(s) have a singular system" % singular)
    if unrepresentable > 0:
        # analysis.py:407 raises ValueError (NaN/inf) or OverflowError (out of int32 range)
        raise ValueError("cannot convert interpolated values to int32 (NaN, inf or out of range)")


def table_device(res: InterpolationResult, out=None):
    """int32 (G, G, P) table of a result on the device (prepare_images_data layout) -- the WHOLE table;
    prepare_images_data and write_table_file go through table_tiles() instead."""
    torch = _torch()
    g = res.axis.size
    p = res.n_pixels
    if out is None:
        out = torch.empty((g, g, p), dtype=torch.int32, device="cuda")
    bad = torch.zeros(2, dtype=torch.int32, device="cuda")
    _table_rows(res, 0, res.rows, out.view(g * g, p), bad)
    _raise_if_bad_table(bad)
    return out


def table_tiles(res: InterpolationResult, tile_bytes=None, with_events=False):
    """Yields (r0, tr, d_tile): rows [r0, r0 + tr) of every image as an int32 (G*G, tr*cols) device
    tensor, produced through two alternating device buffers of at most tile_bytes each; tile t + 1 is
    being computed while the caller consumes tile t.  Numerical failures raise after the last tile.
    with_events: a fourth item, the CUDA event recorded behind the tile's kernels (a consumer on another
    stream waits for that instead of for the next tile, which is already queued behind it)."""
    torch = _torch()
    g = res.axis.size
    gg = g * g
    tile_bytes = TABLE_TILE_BYTES if tile_bytes is None else tile_bytes
    tr_max = int(max(1, min(res.rows, tile_bytes // max(1, gg * 4 * res.cols))))
    if res.mode == "rbf_shared":
        tr_max = res.rows
    if res.banded:
        # whole 16-row pixel tiles per band (the pcg fit's tiles), bounded by the SoA / weights / workspace memory
        free, _total = torch.cuda.mem_get_info()
        per_row = res.cols * res.stack.n * 64                      # SoA + weights + pcg state, bytes per ROI row
        tr_mem = int(max(16, (free // 4) // max(per_row, 1)))
        tr_max = max(16, (min(tr_max, tr_mem) // 16) * 16)
        tr_max = min(tr_max, res.rows)
    # whole 16-byte granules per tile keep the fused kernel's frame rows aligned (its fast path)
    while tr_max < res.rows and (tr_max * res.cols) % 16 != 0:
        tr_max += 1
    tiles = [(r0, min(tr_max, res.rows - r0)) for r0 in range(0, res.rows, tr_max)]
    nbuf = 1 if len(tiles) == 1 else 2
    bufs = [torch.empty(gg * tr_max * res.cols, dtype=torch.int32, device="cuda") for _ in range(nbuf)]
    bad = torch.zeros(2, dtype=torch.int32, device="cuda")

    def issue(t):
        r0, tr = tiles[t]
        view = bufs[t % nbuf][:gg * tr * res.cols].view(gg, tr * res.cols)
        _table_rows(res, r0, tr, view, bad)
        ev = None
        if with_events:
            ev = torch.cuda.Event()
            ev.record()
        return view, ev

    nxt = issue(0)
    for t, (r0, tr) in enumerate(tiles):
        cur, ev = nxt
        if t + 1 < len(tiles):
            nxt = issue(t + 1)             # its buffer was consumed (synchronously) two tiles ago
        if with_events:
            yield r0, tr, cur, ev
        else:
            yield r0, tr, cur
    _raise_if_bad_table(bad)


_PINNED = {}          # (device, bytes) -> two pinned staging tensors, reused across calls


def _stream_images(d_table, consume, chunk_bytes=128 << 20):
    """Device table int32 (n_images, P) -> host, in chunks through two pinned staging buffers:
    the D2H copy of chunk c+1 runs (side stream) while ``consume(host_chunk, i0)`` handles chunk c.
    ``host_chunk`` is an int32 ndarray (count, P) that is only valid during the call."""
    torch = _torch()
    n_images, p = d_table.shape
    per = max(1, min(n_images, chunk_bytes // max(1, p * 4)))
    elems = max(per * p, chunk_bytes // 4)
    key = d_table.device.index
    if key not in _PINNED or _PINNED[key][0].numel() < per * p:
        _PINNED.pop(key, None)
        _PINNED[key] = [torch.empty(elems, dtype=torch.int32, pin_memory=True) for _ in range(2)]
    bufs = [b[:per * p].view(per, p) for b in _PINNED[key]]
    side = torch.cuda.Stream()
    side.wait_stream(torch.cuda.current_stream())
    events = [torch.cuda.Event(), torch.cuda.Event()]
    starts = list(range(0, n_images, per))

    def issue(c):
        i0 = starts[c]
        cnt = min(per, n_images - i0)
        with torch.cuda.stream(side):
            bufs[c & 1][:cnt].copy_(d_table[i0:i0 + cnt], non_blocking=True)
            events[c & 1].record(side)

    issue(0)
    for c, i0 in enumerate(starts):
        if c + 1 < len(starts):
            issue(c + 1)                       # its buffer was consumed two iterations ago
        events[c & 1].synchronize()
        cnt = min(per, n_images - i0)
        consume(bufs[c & 1][:cnt].numpy(), i0)
    torch.cuda.current_stream().wait_stream(side)


def _table_to_host(d_table):
    """(n_images, P) device table -> fresh ndarray, the chunks copied out of the pinned ring by a few
    host threads (a pageable 6.4 GB cudaMemcpy runs at ~2 GB/s; this runs at host-memcpy speed)."""
    from concurrent.futures import ThreadPoolExecutor
    n_images, p = d_table.shape
    out = np.empty((n_images, p), dtype=np.int32)
    threads = max(1, min(8, len(os.sched_getaffinity(0)) if hasattr(os, "sched_getaffinity") else 4))
    with ThreadPoolExecutor(threads) as pool:
        def consume(chunk, i0):
            cnt = chunk.shape[0]
            step = (cnt + threads - 1) // threads
            futs = [pool.submit(np.copyto, out[i0 + a:i0 + min(cnt, a + step)], chunk[a:min(cnt, a + step)])
                    for a in range(0, cnt, step)]
            for f in futs:
                f.result()
        _stream_images(d_table, consume)
    return out


def _table_to_host_tiled(res):
    """The whole (G*G, P) table as a fresh host ndarray, evaluated in row tiles on the device.  Each tile
    leaves through rti_table_tile_to_host (int16 over PCIe when its values fit, a pinned staging ring and
    host threads that write the int32 result); RTI_TABLE_TO_HOST=python keeps the earlier
    route (int32 chunks through a pinned ring, copied out by a Python thread pool) for comparison."""
    from concurrent.futures import ThreadPoolExecutor
    gg = res.axis.size ** 2
    out = np.empty((gg, res.rows, res.cols), dtype=np.int32)
    if gg > 0 and os.environ.get("RTI_TABLE_TO_HOST", "native") != "python":
        torch = _torch()
        p = res.rows * res.cols
        for r0, tr, d_tile, ev in table_tiles(res, with_events=True):
            _capi.call("rti_table_tile_to_host", d_tile.data_ptr(), tr * res.cols, gg, tr * res.cols,
                       out.ctypes.data, p, r0 * res.cols, ev.cuda_event, _stream_ptr(torch))
        return out.reshape(gg, p)
    threads = max(1, min(8, len(os.sched_getaffinity(0)) if hasattr(os, "sched_getaffinity") else 4))
    with ThreadPoolExecutor(threads) as pool:
        for r0, tr, d_tile in table_tiles(res):
            def consume(chunk, i0, r0=r0, tr=tr):
                cnt = chunk.shape[0]
                step = (cnt + threads - 1) // threads
                src = chunk.reshape(cnt, tr, res.cols)
                futs = [pool.submit(np.copyto, out[i0 + a:i0 + min(cnt, a + step), r0:r0 + tr],
                                    src[a:min(cnt, a + step)]) for a in range(0, cnt, step)]
                for f in futs:
                    f.result()
            _stream_images(d_tile, consume)
    return out.reshape(gg, res.rows * res.cols)


def prepare_images_data(data, first_only=False):
    """Prepare images for each light position (analysis.py:375-411).

    :param data: the object returned by interpolate_intensities (or the reference's nested
                 list interpolation_intensities[y][x][ly][lx])
    :return: interpolated_images[ly][lx] -> writable int32 (R, R) ndarray
    """
    if data is None or len(data) <= 0:
        raise Exception("Error preparing images: results are empty")
    print("Preparing images values:")
    torch = _torch()
    range_val = 1 if first_only else _cst.ROI_DIAMETER
    if isinstance(data, InterpolationResult):
        if (data.rows, data.cols) != (range_val, range_val):
            raise IndexError("interpolation result is %dx%d, ROI is %d" % (data.rows, data.cols, range_val))
        g = data.axis.size
        table = _table_to_host_tiled(data).reshape(g, g, range_val, range_val)
    else:
        # plain nested list / array [y][x][ly][lx]: transpose + truncate on the device
        grids = np.ascontiguousarray(
            np.asarray([[np.asarray(data[y][x], dtype=np.float64) for x in range(range_val)]
                        for y in range(range_val)]))
        g = grids.shape[-1]
        p = range_val * range_val
        d_grids = torch.from_numpy(grids).cuda()
        d_table = torch.empty((g, g, p), dtype=torch.int32, device="cuda")
        bad = torch.zeros(1, dtype=torch.int32, device="cuda")
        step = 65535 * 32
        for p0 in range(0, p, step):
            cnt = min(step, p - p0)
            _capi.call("rti_grids_to_table", d_grids.data_ptr() + 8 * p0 * g * g, cnt, g,
                       d_table.data_ptr() + 4 * p0, p, bad.data_ptr(), _stream_ptr(torch))
        if int(bad.item()) > 0:
            raise ValueError("cannot convert interpolated values to int32 (NaN, inf or out of range)")
        table = _table_to_host(d_table.view(g * g, -1)).reshape(g, g, range_val, range_val)
    return [[table[v, u] for u in range(table.shape[1])] for v in range(table.shape[0])]


# ------------------------------------------------------------------------- file I/O
def write_on_file(data, filename, compressed=True):
    """Same files as Utils/utilities.py:48-70: '<name>.pbz2' (bz2 + pickle) or '<name>.pickle'."""
    import bz2
    import pickle
    print("Writing on file {}".format(filename))
    if compressed:
        filename = filename + ".pbz2"
        with bz2.BZ2File(filename, "wb") as f:
            pickle.dump(data, f)
    else:
        filename = filename + ".pickle"
        slab = table_io.is_table_slab(data)          # the nested list of views prepare_images_data returns
        if slab is not None and not table_io.template_ok(slab.shape[2:], slab.dtype):
            slab = None                              # this numpy pickles ndarrays differently: plain pickle.dump
        with open(filename, "wb") as f:
            if slab is not None and slab.nbytes > 0:
                table_io.dump_table(slab, f)         # same object graph, no per-image copies
            else:
                pickle.dump(data, f)
    print("Saved file {}".format(filename))


def read_from_file(filename, compressed=True):
    """Utils/utilities.py:73-101, including its 'Storage file not found!' error."""
    import bz2
    import pickle
    filename = filename + (".pbz2" if compressed else ".pickle")
    if not os.path.isfile(filename):
        raise Exception("Storage file not found!")
    print("Reading from file {}".format(filename))
    if compressed:
        from . import ingest
        return ingest.read_pbz2(filename)        # same object, bzip2 blocks decoded in parallel
    table = table_io.load_table(filename)        # a result table written by this package: mapped, not read
    if table is not None:
        return table
    with open(filename, "rb") as f:
        return pickle.load(f)


def compute(video_name="coin1", from_storage=False, storage_filepath=None, interpolate_PTM=False,
            notification_email=False, debug=False):
    """Main function (analysis.py:414-482), steps 2-4 on the GPU.

    Step 1 from video (from_storage=False) is the reference's acquisition stage and is
    not part of this package.  notification_email is accepted for signature
    compatibility and never sends anything.
    """
    results_frames_filepath = "assets/frames_results_{}".format(video_name)
    results_interpolation_filepath = "assets/interpolation_results_{}".format(video_name)
    print("\nStep 1: Computing frames values")
    if from_storage is True:
        if storage_filepath is not None:
            results_frames_filepath = storage_filepath
        results_frames = read_from_file(results_frames_filepath)
    else:
        raise NotImplementedError(
            "video acquisition (analysis.py:436-460) is outside this package; run the reference's "
            "acquisition once and call compute(from_storage=True)")
    if debug:
        print("Notice: computing in debug mode (first pixel only)")
    print("\nStep 2: Computing pixels intensities")
    data = compute_intensities(results_frames, first_only=debug)
    print("\nStep 3: Computing interpolation")
    results_interpolation = interpolate_intensities(data, interpolate_PTM=interpolate_PTM, first_only=debug)
    print("\nStep 4: Preparing images data")
    if debug:
        results_images = prepare_images_data(results_interpolation, first_only=True)
    else:
        # analysis.py:472-475 as one pipeline: the table is evaluated on the device and goes
        # device -> pinned ring -> file, image by image, without ever being assembled on the host
        results_images = None
        write_table_file(results_interpolation, results_interpolation_filepath)
        if notification_email:
            print("notification_email is ignored (no network side effects)")
    print("\nOK. Computation completed")
    return results_images


def write_table_file(res, filename):
    """prepare_images_data + ut.write_on_file(.., compressed=False) (analysis.py:472-475) fused: writes
    '<filename>.pickle' with the reference's object graph (list[G] of list[G] of int32 (R, R) arrays)."""
    if not isinstance(res, InterpolationResult):
        return write_on_file(prepare_images_data(res), filename, compressed=False)
    if (res.rows, res.cols) != (_cst.ROI_DIAMETER, _cst.ROI_DIAMETER):
        raise IndexError("interpolation result is %dx%d, ROI is %d" % (res.rows, res.cols, _cst.ROI_DIAMETER))
    print("Preparing images values:")
    g = res.axis.size
    if not table_io.template_ok((res.rows, res.cols), np.int32):
        # this numpy pickles ndarrays in a way the streaming writers do not know: plain pickle.dump
        return write_on_file(prepare_images_data(res), filename, compressed=False)
    print("Writing on file {}".format(filename))
    filename = filename + ".pickle"
    from concurrent.futures import ThreadPoolExecutor
    writer = table_io.TableTileWriter(g, g, (res.rows, res.cols), np.int32, filename)
    threads = max(1, min(8, len(os.sched_getaffinity(0)) if hasattr(os, "sched_getaffinity") else 4))
    try:
        if g == 0:
            pass                                     # an empty light grid: the file is the empty list already
        elif os.environ.get("RTI_TABLE_TO_HOST", "native") != "python":
            # row tiles of the ROI: device tile -> (int16 over PCIe when it fits) -> pinned ring -> host threads
            # that pwrite each image's rows at its fixed offset (rti_table_tile_to_file); the evaluation of the
            # next tile runs meanwhile
            torch = _torch()
            offsets = np.ascontiguousarray(writer.offsets, dtype=np.int64)
            for r0, tr, d_tile, ev in table_tiles(res, with_events=True):
                _capi.call("rti_table_tile_to_file", d_tile.data_ptr(), tr * res.cols, g * g, tr * res.cols,
                           writer.fd, offsets.ctypes.data, r0 * res.cols, ev.cuda_event, _stream_ptr(torch))
        else:
            # the earlier route: int32 chunks through a pinned ring, pwrite from a Python thread pool
            with ThreadPoolExecutor(threads) as pool:
                for r0, tr, d_tile in table_tiles(res):
                    def consume(chunk, i0, r0=r0):
                        cnt = chunk.shape[0]
                        step = (cnt + threads - 1) // threads
                        futs = [pool.submit(writer.write_rows, i0 + a, chunk[a:min(cnt, a + step)], r0)
                                for a in range(0, cnt, step)]
                        for f in futs:
                            f.result()
                    _stream_images(d_tile, consume)
    finally:
        writer.close()
    print("Saved file {}".format(filename))


def install(analysis_module, constants_module=None, fast_writer=True, fast_reader=True):
    """Drop-in: rebind the reference module's hot functions to this implementation.

    analysis.compute() resolves compute_intensities / interpolate_intensities /
    prepare_images_data through its module globals at call time (analysis.py:466-472), so
    after ``install(analysis)`` the unmodified reference ``compute()`` runs on the GPU.
    The tunables are then read from the reference's own ``constants`` module.
    """
    global _cst
    if constants_module is None:
        constants_module = getattr(analysis_module, "cst", None) or sys.modules.get("constants")
    if constants_module is not None:
        _cst = constants_module
    analysis_module.compute_intensities = compute_intensities
    analysis_module.interpolate_intensities = interpolate_intensities
    analysis_module.prepare_images_data = prepare_images_data
    # analysis.compute() stores the table with ut.write_on_file(.., compressed=False) (analysis.py:475):
    # route that one call shape through the streaming writer; everything else goes to the original
    ut = getattr(analysis_module, "ut", None)
    if fast_writer and ut is not None and hasattr(ut, "write_on_file") \
            and not getattr(ut.write_on_file, "_rti_b200", False):
        original = ut.write_on_file

        def write_on_file_fast(data, filename, compressed=True):
            if compressed or table_io.is_table_slab(data) is None:
                return original(data, filename, compressed=compressed)
            return write_on_file(data, filename, compressed=False)

        write_on_file_fast._rti_b200 = True
        write_on_file_fast._original = original
        ut.write_on_file = write_on_file_fast
    # analysis.compute(from_storage=True) loads the frame list with ut.read_from_file (analysis.py:434)
    if fast_reader and ut is not None and hasattr(ut, "read_from_file") \
            and not getattr(ut.read_from_file, "_rti_b200", False):
        original_reader = ut.read_from_file

        def read_from_file_fast(filename, compressed=True):
            # compressed: parallel bzip2 blocks; uncompressed (interactive_relighting.py:95): a table
            # written by this package is memory-mapped copy-on-write, anything else is unpickled as before
            return read_from_file(filename, compressed=compressed)

        read_from_file_fast._rti_b200 = True
        read_from_file_fast._original = original_reader
        ut.read_from_file = read_from_file_fast
    return analysis_module
