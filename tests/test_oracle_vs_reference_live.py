"""Live differential pin of the CPU oracle: property tests (hypothesis) that run the UNMODIFIED
reference functions imported from /root/reference on random small inputs and hold the oracle to them.
Runs only where the reference is mounted (the build container; `-m "not gpu"` there); the GPU box
has no reference and uses the committed goldens instead."""
import os
import sys
import types

import numpy as np
import pytest

import rti_oracle as orc

REF = os.environ.get("RTI_REFERENCE_DIR", "/root/reference")
pytestmark = pytest.mark.skipif(not os.path.isfile(os.path.join(REF, "analysis.py")),
                                reason="reference not mounted on this machine")

hypothesis = pytest.importorskip("hypothesis")
from hypothesis import HealthCheck, given, settings  # noqa: E402
from hypothesis import strategies as st  # noqa: E402

COMMON = dict(deadline=None, suppress_health_check=list(HealthCheck), derandomize=True)


@pytest.fixture(scope="module")
def ref():
    """(analysis, constants, utilities) of the reference, matplotlib stubbed (analysis.py:14 imports it
    for the first_only debug plots only)."""
    mpl = types.ModuleType("matplotlib")
    plt = types.ModuleType("matplotlib.pyplot")
    mpl.pyplot = plt
    sys.modules.setdefault("matplotlib", mpl)
    sys.modules.setdefault("matplotlib.pyplot", plt)
    sys.path.insert(0, REF)
    cwd = os.getcwd()
    os.chdir(REF)
    try:
        import analysis
        import constants
        from Utils import utilities
    finally:
        os.chdir(cwd)
    old = (constants.ROI_DIAMETER, constants.INTERPOLATION_PARAM)
    yield analysis, constants, utilities
    constants.ROI_DIAMETER, constants.INTERPOLATION_PARAM = old


def _frames(rng, r, n):
    cams = np.column_stack([rng.uniform(-1.5 * r, 2.5 * r, n), rng.uniform(-1.5 * r, 2.5 * r, n),
                            rng.uniform(0.75 * r, 2.25 * r + 1, n)])
    return [(rng.integers(0, 256, (r, r), dtype=np.uint8), cams[i]) for i in range(n)]


@settings(max_examples=25, **COMMON)
@given(r=st.integers(1, 7), n=st.integers(1, 30), seed=st.integers(0, 2 ** 31 - 1))
def test_compute_intensities_bit_exact(ref, r, n, seed):
    analysis, cst, _ = ref
    cst.ROI_DIAMETER = r
    recs = _frames(np.random.default_rng(seed), r, n)
    lx, ly, inten = analysis.compute_intensities(recs)
    o_lx, o_ly, o_in = orc.gather_soa(np.stack([f[0] for f in recs]), np.stack([f[1] for f in recs]))
    assert np.array_equal(o_in, inten) and inten.dtype == o_in.dtype
    assert np.array_equal(o_lx, lx) and np.array_equal(o_ly, ly)


@settings(max_examples=12, **COMMON)
@given(n=st.integers(6, 80), seed=st.integers(0, 2 ** 31 - 1))
def test_interpolate_ptm_bit_exact(ref, n, seed):
    """_interpolate_PTM (analysis.py:263-317): oracle grids bit for bit (libm pow reproduced), and the
    correctly rounded float32 products of the CUDA path within 1e-6."""
    analysis, cst, _ = ref
    rng = np.random.default_rng(seed)
    rr, th = np.sqrt(rng.uniform(0, 0.95, n)), rng.uniform(0, 2 * np.pi, n)
    lx, ly = (rr * np.cos(th)).astype(np.float32), (rr * np.sin(th)).astype(np.float32)
    val = rng.integers(0, 256, n).astype(np.int32)
    ax = orc.light_axis(cst.INTERPOLATION_PARAM)
    want = analysis._interpolate_PTM(lx, ly, ax, val)
    got = orc.ptm_grid(orc.ptm_fit(lx, ly, val, libm_pow=True), ax)
    assert np.array_equal(got, want)
    plain = orc.ptm_grid(orc.ptm_fit(lx, ly, val), ax)
    assert np.max(np.abs(plain - want)) <= 1e-6 * max(1.0, np.max(np.abs(want)))


@settings(max_examples=12, **COMMON)
@given(n=st.integers(3, 90), seed=st.integers(0, 2 ** 31 - 1))
def test_interpolate_rbf_matches_scipy_path(ref, n, seed):
    """_interpolate_RBF (analysis.py:249-260 -> scipy Rbf 'linear'): 1e-9 of the largest value (numpy vs
    scipy solve drivers differ at round-off)."""
    analysis, cst, _ = ref
    rng = np.random.default_rng(seed)
    rr, th = np.sqrt(rng.uniform(0, 0.95, n)), rng.uniform(0, 2 * np.pi, n)
    lx, ly = (rr * np.cos(th)).astype(np.float32), (rr * np.sin(th)).astype(np.float32)
    if len({(a, b) for a, b in zip(lx.tolist(), ly.tolist())}) < n:
        return                                      # duplicate nodes: singular in both (separate test)
    val = rng.integers(0, 256, n).astype(np.int32)
    ax = orc.light_axis(cst.INTERPOLATION_PARAM)
    yi, xi = np.meshgrid(ax, ax, indexing="ij")
    want = analysis._interpolate_RBF(lx, ly, xi, yi, val)
    got = orc.rbf_grid(lx, ly, orc.rbf_fit(lx, ly, val), ax)
    assert got.shape == want.shape
    assert np.max(np.abs(got - want)) <= 1e-9 * max(1.0, np.max(np.abs(want)))


@settings(max_examples=6, **COMMON)
@given(r=st.integers(1, 3), n=st.integers(6, 20), ptm=st.booleans(), seed=st.integers(0, 2 ** 31 - 1))
def test_whole_path_table(ref, r, n, ptm, seed):
    """compute_intensities -> interpolate_intensities -> prepare_images_data on a tiny ROI with a coarse
    light grid (INTERPOLATION_PARAM = 0.25): the int32 table, including its truncation."""
    analysis, cst, _ = ref
    cst.ROI_DIAMETER, cst.INTERPOLATION_PARAM = r, 0.25
    try:
        recs = _frames(np.random.default_rng(seed), r, n)
        data = analysis.compute_intensities(recs)
        interp = analysis.interpolate_intensities(data, interpolate_PTM=ptm)
        want = np.array(analysis.prepare_images_data(interp))               # (g, g, r, r) int32
    finally:
        cst.INTERPOLATION_PARAM = 0.02
    grids = orc.interpolate(data[0], data[1], data[2], ptm=ptm, step=0.25, **({"libm_pow": True} if ptm else {}))
    got = orc.table_from_grids(grids)
    assert got.dtype == want.dtype and got.shape == want.shape
    diff = got != want
    if ptm:
        assert not diff.any()
    else:
        ref_f = np.array([[interp[y][x] for x in range(r)] for y in range(r)]).transpose(2, 3, 0, 1)
        assert np.all(np.abs(ref_f[diff] - np.rint(ref_f[diff])) < 1e-8)


@settings(max_examples=40, **COMMON)
@given(w=st.integers(40, 130), h=st.integers(40, 100), diam=st.integers(2, 60),
       cx=st.integers(0, 129), cy=st.integers(0, 99), ex=st.integers(0, 30), ey=st.integers(0, 30),
       seed=st.integers(0, 2 ** 31 - 1))
def test_get_roi_hsv(ref, w, h, diam, cx, cy, ex, ey, seed):
    """ut.get_ROI(img, corners, hsv=True) (Utils/utilities.py:302-342) for windows that hit every
    clamping branch."""
    _, cst, ut = ref
    cst.ROI_DIAMETER = diam
    cx, cy = min(cx, w - 1), min(cy, h - 1)
    corners = [(cx, cy), (min(cx + ex, w - 1), cy), (min(cx + ex, w - 1), min(cy + ey, h - 1)), (cx, min(cy + ey, h - 1))]
    img = np.random.default_rng(seed).integers(0, 256, (h, w, 3), dtype=np.uint8)
    win = orc.roi_bounds(img.shape, corners, diam)
    if win[1] <= win[0] or win[3] <= win[2]:
        return                                      # empty crop: cv2.cvtColor raises in the reference
    want = ut.get_ROI(img, corners, hsv=True)
    got = orc.get_roi_hsv(img, corners, diam)
    assert got.shape == want.shape and np.array_equal(got, want)


@settings(max_examples=60, **COMMON)
@given(x=st.integers(0, 400), y=st.integers(0, 400), hh=st.sampled_from([200, 400, 401]))
def test_cursor_to_light_index(ref, x, y, hh):
    _, _, ut = ref
    x, y = min(x, hh), min(y, hh)
    lx, ly = ut.draw_light_roi_position(x, y, (hh, hh), to_light_vector=True)
    assert orc.cursor_to_light(x, y, hh, hh) == (lx, ly)
    assert orc.light_to_index(lx) == round((1 + lx) / 2 * 100)
