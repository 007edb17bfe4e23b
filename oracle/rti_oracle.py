"""CPU oracle for the per-pixel reflectance-fitting path of Smartphone-based-RTI.

TEST INFRASTRUCTURE ONLY.  Nothing under ``smartphone-based-rti_b200/`` (the
product) may import this module; it is used by ``tests/``, by
``__graft_entry__.smoke()`` and by the ``cpu_baseline`` / ``--impl reference``
legs of ``bench.py`` as the checker / timed CPU arm, never as a fallback.

It is a numpy (fp64) restatement of the reference's algorithm, each function
citing the reference lines it follows (paths relative to the upstream repo):

    analysis.py:196-246   compute_intensities      -> gather_soa
    analysis.py:263-317   _interpolate_PTM         -> ptm_design / ptm_fit / ptm_grid
    analysis.py:249-260   _interpolate_RBF         -> rbf_fit / rbf_grid
      (scipy 1.18.1 interpolate/_rbf.py Rbf.__init__/A/__call__, function='linear';
       third-party, not vendored by the reference, which pins no version)
    analysis.py:321-372   interpolate_intensities  -> light_axis / interpolate
    analysis.py:375-411   prepare_images_data      -> table_from_grids
    interactive_relighting.py:11-39 + Utils/utilities.py:368-376
                                                   -> cursor_to_light / light_to_index /
                                                      clamp_values / hsv2bgr_u8
    Utils/utilities.py:272-342 get_corners_center / get_ROI(hsv=True),
    FeatureMatcher.py:183-184                      -> corners_center / roi_bounds /
                                                      bgr2hsv_u8 / get_roi_hsv / roi_intensities

Parity pin: the reference has no tests or golden vectors of its own
(SURVEY.md section 4).  This oracle is pinned instead against outputs of the
reference itself, produced in the build container by
``tests/golden/make_golden.py`` (which imports the unmodified reference) and
committed under ``tests/golden/``; ``tests/test_oracle_golden.py`` checks it.
"""
from __future__ import annotations

import numpy as np

__all__ = [
    "light_axis", "gather_soa", "ptm_design", "ptm_fit", "ptm_grid", "rbf_matrix",
    "rbf_fit", "rbf_grid", "interpolate", "table_from_grids", "cursor_to_light",
    "light_to_index", "clamp_values", "hsv2bgr_u8", "relight_from_table",
    "corners_center", "roi_bounds", "bgr2hsv_u8", "get_roi_hsv", "roi_intensities",
]


# --------------------------------------------------------------------------- grid
def light_axis(step: float = 0.02) -> np.ndarray:
    """One axis of the light-direction grid (analysis.py:345-347, 390-393).

    ``np.mgrid[-1:1:step]`` rounded to 2 decimals; for step 0.02 this is the
    100 values -1.00 ... 0.98.  Both grid axes use the same values (xi[0]).
    """
    return np.around(np.mgrid[-1:1:step], decimals=2)


# ------------------------------------------------------------------------- gather
def gather_soa(frames, cams, rows: int | None = None, cols: int | None = None, row0: int = 0,
               col0: int = 0):
    """Light vectors and intensities per pixel and frame (analysis.py:217-232).

    frames: (N, H, W) uint8, cams: (N, 3) float64 (camera position per frame).
    For pixel p=(x, y, 0) and frame i: l = (cam_i - p) / ||cam_i - p|| in fp64;
    l[0], l[1] are stored as float32, the intensity as int32, all with the frame
    index fastest: arrays of shape (rows, cols, N).

    row0 / col0: `frames` is the window of the ROI that starts at pixel (x, y) = (col0, row0);
    the reference walks the whole ROI, a window evaluates the same formula for its pixels.
    """
    frames = np.asarray(frames)
    cams = np.asarray(cams, dtype=np.float64)
    n = frames.shape[0]
    rows = frames.shape[1] if rows is None else rows
    cols = frames.shape[2] if cols is None else cols
    ys = np.arange(row0, row0 + rows, dtype=np.float64)[:, None, None]
    xs = np.arange(col0, col0 + cols, dtype=np.float64)[None, :, None]
    dx = cams[None, None, :, 0] - xs                      # (1, cols, N)
    dy = cams[None, None, :, 1] - ys                      # (rows, 1, N)
    dz = cams[None, None, :, 2]
    norm = np.sqrt(dx * dx + dy * dy + dz * dz)           # np.linalg.norm, analysis.py:229
    lx = (dx / norm).astype(np.float32)                   # stored into a float32 array, :230
    ly = (dy / norm).astype(np.float32)                   # :231
    inten = np.ascontiguousarray(
        np.transpose(frames[:, :rows, :cols], (1, 2, 0))).astype(np.int32)   # :232
    assert lx.shape == (rows, cols, n)
    return lx, ly, inten


# ---------------------------------------------------------------------------- PTM
def _square_like_reference(a: np.ndarray) -> np.ndarray:
    """``lu ** 2`` as the reference evaluates it (analysis.py:285): a numpy
    float32 SCALAR raised to the Python int 2 goes through libm ``powf``, which
    is within 1 ulp but not correctly rounded -- about 0.07 % of values differ
    from the float32 product ``lu * lu`` by one ulp (measured here, numpy 2.3.5 /
    glibc).  Slow Python loop; only used to pin the oracle bit-for-bit."""
    flat = np.asarray(a, dtype=np.float32).ravel()
    return np.array([v ** 2 for v in flat], dtype=np.float32).reshape(np.shape(a))


def ptm_design(lx, ly, libm_pow: bool = False) -> np.ndarray:
    """N x 6 design matrix (analysis.py:282-290).

    lx, ly are float32; the three products are formed in float32 (numpy float32
    scalars stay float32 under ``**2`` and ``*``) and only then promoted to fp64
    by ``np.array`` because the row also holds the Python float ``1.``.

    libm_pow=False forms the squares as correctly rounded float32 products (what
    the CUDA path computes); libm_pow=True reproduces the reference's scalar
    ``** 2`` bit for bit (see _square_like_reference).  The two differ by one
    float32 ulp in ~0.07 % of entries, i.e. ~4e-8 relative in the coefficients.
    """
    lu = np.asarray(lx, dtype=np.float32)
    lv = np.asarray(ly, dtype=np.float32)
    m = np.empty(lu.shape + (6,), dtype=np.float64)
    m[..., 0] = _square_like_reference(lu) if libm_pow else lu * lu
    m[..., 1] = _square_like_reference(lv) if libm_pow else lv * lv
    m[..., 2] = lu * lv
    m[..., 3] = lu
    m[..., 4] = lv
    m[..., 5] = 1.0
    return m


def ptm_fit(lx, ly, values, libm_pow: bool = False) -> np.ndarray:
    """Six biquadratic coefficients of one pixel (analysis.py:295-298).

    Full SVD of the design matrix, ``a = V^T.T @ ((U.T @ L)[:6] / s)``: the
    minimum-norm least-squares solution without a rank cut-off.  N < 6 raises
    ValueError exactly like the reference (shapes not aligned in the last dot).
    """
    a_mat = ptm_design(lx, ly, libm_pow)
    rhs = np.asarray(values)
    u, s, vt = np.linalg.svd(a_mat)
    c = np.dot(u.T, rhs)
    w = np.divide(c[:len(s)], s)
    return np.dot(vt.T, w)


def ptm_fit_batch(lx, ly, values) -> np.ndarray:
    """ptm_fit over leading axes (same algorithm, reduced SVD, batched)."""
    a_mat = ptm_design(lx, ly)                                    # (..., N, 6)
    rhs = np.asarray(values, dtype=np.float64)[..., None]          # (..., N, 1)
    if a_mat.shape[-2] < 6:
        raise ValueError("PTM fit needs at least 6 light samples")
    u, s, vt = np.linalg.svd(a_mat, full_matrices=False)
    c = np.matmul(np.swapaxes(u, -1, -2), rhs)[..., 0]            # (..., 6)
    return np.matmul(np.swapaxes(vt, -1, -2), (c / s)[..., None])[..., 0]


def ptm_grid(coef, axis) -> np.ndarray:
    """Evaluate the biquadratic on the grid (analysis.py:301-315).

    Returns (G, G) fp64 indexed [v(ly)][u(lx)]; the sum is formed left to right
    exactly as the reference does so the fp64 result is bit-identical.
    """
    a = np.asarray(coef, dtype=np.float64)
    ax = np.asarray(axis, dtype=np.float64)
    lu = ax[None, :]
    lv = ax[:, None]
    out = a[..., 0, None, None] * (lu ** 2)
    out = out + a[..., 1, None, None] * (lv ** 2)
    out = out + a[..., 2, None, None] * (lu * lv)
    out = out + a[..., 3, None, None] * lu
    out = out + a[..., 4, None, None] * lv
    out = out + a[..., 5, None, None]
    return out


# ---------------------------------------------------------------------------- RBF
def rbf_matrix(lx, ly) -> np.ndarray:
    """Kernel matrix of scipy's legacy ``Rbf(function='linear')``.

    A[j, k] = ||n_j - n_k||_2 with nodes cast to fp64 (scipy _rbf.py Rbf.A:
    squareform(pdist(xi.T)); 'linear' is phi(r) = r; smooth = 0).
    """
    x = np.asarray(lx, dtype=np.float64)
    y = np.asarray(ly, dtype=np.float64)
    ddx = x[..., :, None] - x[..., None, :]
    ddy = y[..., :, None] - y[..., None, :]
    return np.sqrt(ddx * ddx + ddy * ddy)


def rbf_fit(lx, ly, values) -> np.ndarray:
    """RBF weights of one pixel: solve(A, d) (scipy _rbf.py Rbf.__init__).

    Duplicate nodes make A singular and raise numpy's LinAlgError like
    scipy.linalg.solve does in the reference.
    """
    a_mat = rbf_matrix(lx, ly)
    d = np.asarray(values, dtype=np.float64)
    return np.linalg.solve(a_mat, d[..., None])[..., 0]


def rbf_grid(lx, ly, weights, axis) -> np.ndarray:
    """Evaluate the interpolant on the grid (scipy _rbf.py Rbf.__call__).

    out[v, u] = sum_k ||(axis[u], axis[v]) - n_k|| * w_k, (G, G) fp64 indexed
    [ly][lx] (analysis.py:345, 361-363: x_fine = xi, y_fine = yi of the mgrid).
    """
    x = np.asarray(lx, dtype=np.float64)
    y = np.asarray(ly, dtype=np.float64)
    w = np.asarray(weights, dtype=np.float64)
    ax = np.asarray(axis, dtype=np.float64)
    gx = ax[None, :, None]
    gy = ax[:, None, None]
    r = np.sqrt((gx - x[None, None, :]) ** 2 + (gy - y[None, None, :]) ** 2)
    return r @ w


# ------------------------------------------------------------------------- driver
def interpolate(lx, ly, inten, ptm: bool, step: float = 0.02,
                libm_pow: bool = False) -> np.ndarray:
    """interpolate_intensities as one array (analysis.py:321-372).

    Returns (rows, cols, G, G) fp64: [y][x][ly][lx].
    """
    ax = light_axis(step)
    rows, cols, _ = lx.shape
    out = np.empty((rows, cols, ax.size, ax.size), dtype=np.float64)
    for y in range(rows):
        for x in range(cols):
            if ptm:
                out[y, x] = ptm_grid(ptm_fit(lx[y, x], ly[y, x], inten[y, x], libm_pow), ax)
            else:
                out[y, x] = rbf_grid(lx[y, x], ly[y, x],
                                     rbf_fit(lx[y, x], ly[y, x], inten[y, x]), ax)
    return out


def table_from_grids(grids) -> np.ndarray:
    """prepare_images_data (analysis.py:400-409).

    (rows, cols, G, G) fp64 -> (G, G, rows, cols) int32: table[ly][lx][y][x].
    Element assignment of a Python float into an int32 array truncates toward
    zero and raises on NaN / inf / out-of-range values; so does this.
    """
    g = np.asarray(grids, dtype=np.float64)
    if not np.all(np.isfinite(g)):
        raise ValueError("cannot convert float NaN/inf to integer")
    if np.any(np.abs(g) >= 2147483648.0):
        raise OverflowError("Python integer out of bounds for int32")
    return np.ascontiguousarray(np.transpose(np.trunc(g), (2, 3, 0, 1))).astype(np.int32)


# ---------------------------------------------------------------------- relighting
def cursor_to_light(x: int, y: int, h: int, w: int):
    """Mouse position -> (lx, ly) (Utils/utilities.py:368-376)."""
    lx = round(2 * (x / w) - 1, 2)
    ly = round(2 * (y / h) - 1, 2)
    if lx >= 0.99:
        lx = 0.98
    if ly >= 0.99:
        ly = 0.98
    return lx, ly


def light_to_index(l: float) -> int:
    """(lx or ly) -> table index (interactive_relighting.py:25-26); Python
    round() is round-half-even."""
    return round((1 + l) / 2 * 100)


def clamp_values(img) -> np.ndarray:
    """values[values > 255] = 255; values[values <= 0] = 0
    (interactive_relighting.py:35-36), out of place here."""
    return np.clip(np.asarray(img), 0, 255)


def hsv2bgr_u8(hsv, simd_lanes: int = 32) -> np.ndarray:
    """8-bit HSV -> BGR with OpenCV's cv2.COLOR_HSV2BGR semantics
    (interactive_relighting.py:38): H in [0,180), S and V in [0,255].

    OpenCV is third-party to the reference (no version pinned; 4.13.0 in the
    build container).  This restates what its 8-bit kernel computes, established
    by exhaustive comparison over all 180*256*256 inputs, once with every pixel in
    the vector body and once with every pixel in the scalar tail
    (tests/golden/make_golden.py, 0 mismatches each): h' = H*(6/180), sector =
    trunc(h'), f = h' - sector, s = S/255, v = V/255 in float32,
        p = v*(1 - s), q = v*fma(-s, f, 1), t = v*fma(-s, 1 - f, 1),
    channel = TRUNCATE(x * 255) in the vector body, ROUND-half-even(x * 255) in
    the scalar tail.  cvtColor converts image rows one at a time, so for a
    (rows, cols, 3) image the tail is the last cols % simd_lanes pixels OF EACH ROW
    (columns >= cols - cols % simd_lanes); simd_lanes = 32 is the uint8 lane count
    of the AVX2 build that runs here and on the GPU box's host (16 on an SSE-only
    host).  Inputs without a row structure (ndim != 3) are treated as vector body.
    """
    hsv = np.asarray(hsv, dtype=np.uint8)
    f32 = np.float32
    h = hsv[..., 0].astype(f32) * f32(6.0 / 180.0)
    s = hsv[..., 1].astype(f32) * f32(1.0 / 255.0)
    v = hsv[..., 2].astype(f32) * f32(1.0 / 255.0)
    sector = np.trunc(h).astype(np.int32)
    f = h - sector.astype(f32)
    sector = sector % 6
    one = f32(1.0)

    def fma(a, b, c):          # single-rounded a*b+c (exact in fp64 for these magnitudes)
        return (a.astype(np.float64) * b.astype(np.float64) + np.float64(c)).astype(f32)

    t1 = v * (one - s)
    t2 = v * fma(-s, f, one)
    t3 = v * fma(-s, one - f, one)
    tab = np.stack([v, t1, t2, t3], axis=-1)
    sel = np.array([[1, 3, 0], [1, 0, 2], [3, 0, 1], [0, 2, 1], [0, 1, 3], [2, 1, 0]])
    bgr = np.take_along_axis(tab, sel[sector], axis=-1)   # b, g, r table slots
    scaled = bgr * f32(255.0)
    out = np.trunc(scaled)
    if hsv.ndim == 3 and simd_lanes > 0:
        cols = hsv.shape[1]
        tail = cols - cols % simd_lanes
        out[:, tail:] = np.rint(scaled[:, tail:])          # cvRound: round half to even
    return np.clip(out, 0, 255).astype(np.uint8)


def relight_from_table(table, hsv_roi, lx: float, ly: float) -> np.ndarray:
    """relighting_event (interactive_relighting.py:25-38) without the window:
    look the image up, clamp, write it into V, convert HSV -> BGR."""
    values = clamp_values(table[light_to_index(ly)][light_to_index(lx)])
    img = np.array(hsv_roi, dtype=np.uint8, copy=True)
    img[:, :, 2] = values
    return hsv2bgr_u8(img)


# --------------------------------------------------------------- ROI extraction
def corners_center(corners, height: int, width: int):
    """get_corners_center (Utils/utilities.py:272-299): centre of the corners'
    bounding box, Python round() (half to even) of max - (max - min)/2."""
    max_x, min_x, max_y, min_y = 0, width, 0, height
    for x, y in corners:
        max_x, min_x = max(max_x, x), min(min_x, x)
        max_y, min_y = max(max_y, y), min(min_y, y)
    return round(max_x - (max_x - min_x) / 2), round(max_y - (max_y - min_y) / 2)


def roi_bounds(shape, corners, roi_diameter: int):
    """The crop window of get_ROI (Utils/utilities.py:311-327), quirks included:
    a negative lower bound becomes 1 (not 0), an upper bound past the edge becomes
    size - 1.  Returns (x_min, x_max, y_min, y_max), upper bounds exclusive."""
    radius = round(roi_diameter / 2)
    h, w = shape[0], shape[1]
    xc, yc = corners_center(np.array(corners), h, w)
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


def bgr2hsv_u8(bgr) -> np.ndarray:
    """8-bit BGR -> HSV with cv2.COLOR_BGR2HSV semantics (Utils/utilities.py:339,
    H in [0,180)).  OpenCV is third-party to the reference (no version pinned; 4.13.0
    in the build container); this restates its integer kernel
        v = max, diff = max - min, s = (diff*sdiv[v] + 2^11) >> 12,
        h = (v==r ? g-b : v==g ? b-r+2diff : r-g+4diff), h = (h*hdiv[diff] + 2^11) >> 12,
        h += 180 if h < 0;  sdiv[i] = rint(255*4096/i), hdiv[i] = rint(180*4096/(6i))
    and was compared with cv2 over all 2^24 colours (tests/golden/make_golden.py)."""
    bgr = np.asarray(bgr, dtype=np.uint8)
    b, g, r = (bgr[..., k].astype(np.int64) for k in range(3))
    v = np.maximum(np.maximum(b, g), r)
    diff = v - np.minimum(np.minimum(b, g), r)
    i = np.arange(1, 256, dtype=np.float64)
    sdiv = np.zeros(256, np.int64)
    hdiv = np.zeros(256, np.int64)
    sdiv[1:] = np.rint((255 << 12) / i).astype(np.int64)
    hdiv[1:] = np.rint((180 << 12) / (6.0 * i)).astype(np.int64)
    s = (diff * sdiv[v] + (1 << 11)) >> 12
    h = np.where(v == r, g - b, np.where(v == g, b - r + 2 * diff, r - g + 4 * diff))
    h = (h * hdiv[diff] + (1 << 11)) >> 12
    h = h + np.where(h < 0, 180, 0)
    return np.stack([h, s, v], axis=-1).astype(np.uint8)


def get_roi_hsv(img_bgr, corners, roi_diameter: int) -> np.ndarray:
    """get_ROI(static_img, points, hsv=True) (Utils/utilities.py:302-340)."""
    x_min, x_max, y_min, y_max = roi_bounds(img_bgr.shape, corners, roi_diameter)
    return bgr2hsv_u8(np.asarray(img_bgr)[y_min:y_max, x_min:x_max])


def roi_intensities(img_bgr, corners, roi_diameter: int) -> np.ndarray:
    """The frame record's first member: V of the HSV ROI (FeatureMatcher.py:183-184)."""
    return get_roi_hsv(img_bgr, corners, roi_diameter)[:, :, 2]
