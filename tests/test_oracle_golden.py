"""The CPU oracle against outputs of the unmodified reference (tests/golden/*.npz, made by
tests/golden/make_golden.py in the build container)."""
import numpy as np
import pytest

import rti_oracle as orc


def test_light_axis():
    ax = orc.light_axis()
    assert ax.shape == (100,) and ax[0] == -1.0 and ax[50] == 0.0 and ax[-1] == 0.98
    assert np.allclose(np.diff(ax), 0.02)


def test_gather_bit_exact(ref_small):
    lx, ly, inten = orc.gather_soa(ref_small["frames"], ref_small["cams"])
    assert lx.dtype == np.float32 and inten.dtype == np.int32
    assert np.array_equal(inten, ref_small["inten"])
    assert np.array_equal(lx, ref_small["lx"])
    assert np.array_equal(ly, ref_small["ly"])


def test_ptm_grids_bit_exact_with_libm_pow(ref_small):
    ax = orc.light_axis()
    for k, (y, x) in enumerate(ref_small["keep"]):
        a = orc.ptm_fit(ref_small["lx"][y, x], ref_small["ly"][y, x], ref_small["inten"][y, x], libm_pow=True)
        assert np.array_equal(orc.ptm_grid(a, ax), ref_small["ptm_grids"][k])


def test_ptm_product_rounding_is_below_tolerance(ref_small):
    """Correctly rounded float32 squares (what the GPU computes) vs the reference's libm powf."""
    ax = orc.light_axis()
    for k, (y, x) in enumerate(ref_small["keep"]):
        a = orc.ptm_fit(ref_small["lx"][y, x], ref_small["ly"][y, x], ref_small["inten"][y, x])
        g = orc.ptm_grid(a, ax)
        ref = ref_small["ptm_grids"][k]
        assert np.max(np.abs(g - ref)) <= 1e-6 * np.max(np.abs(ref))


def test_ptm_batch_matches_per_pixel(ref_small):
    lx, ly, inten = ref_small["lx"], ref_small["ly"], ref_small["inten"]
    batch = orc.ptm_fit_batch(lx, ly, inten)
    for y, x in ref_small["keep"]:
        single = orc.ptm_fit(lx[y, x], ly[y, x], inten[y, x])
        assert np.allclose(batch[y, x], single, rtol=1e-10, atol=1e-10)


def test_ptm_table_exact(ref_small):
    grids = orc.interpolate(ref_small["lx"], ref_small["ly"], ref_small["inten"], ptm=True, libm_pow=True)
    assert np.array_equal(orc.table_from_grids(grids), ref_small["table_ptm"])


def test_rbf_grids_and_table(ref_small):
    grids = orc.interpolate(ref_small["lx"], ref_small["ly"], ref_small["inten"], ptm=False)
    for k, (y, x) in enumerate(ref_small["keep"]):
        ref = ref_small["rbf_grids"][k]
        assert np.max(np.abs(grids[y, x] - ref)) <= 1e-10 * np.max(np.abs(ref))
    table = orc.table_from_grids(grids)
    diff = table != ref_small["table_rbf"]
    # a flip is only legitimate where the fp64 value sits on an integer boundary
    frac = np.abs(np.transpose(grids, (2, 3, 0, 1)) - np.rint(np.transpose(grids, (2, 3, 0, 1))))
    assert np.all(frac[diff] < 1e-9)


def test_rbf_n96(ref_rbf_n96):
    g = ref_rbf_n96
    for k, (y, x) in enumerate(g["pix"]):
        w = orc.rbf_fit(g["lx"][y, x], g["ly"][y, x], g["inten"][y, x])
        out = orc.rbf_grid(g["lx"][y, x], g["ly"][y, x], w, g["axis"])
        assert np.max(np.abs(out - g["rbf_grids"][k])) <= 1e-9 * np.max(np.abs(g["rbf_grids"][k]))
        # interpolant reproduces the node values (SURVEY 8c KAT 3)
        a_mat = orc.rbf_matrix(g["lx"][y, x], g["ly"][y, x])
        assert np.allclose(a_mat @ w, g["inten"][y, x], atol=1e-8)


def test_ptm_analytic_kat():
    """Intensities generated from a known biquadratic are recovered exactly."""
    rng = np.random.default_rng(5)
    lx = rng.uniform(-0.9, 0.9, 80).astype(np.float32)
    ly = rng.uniform(-0.9, 0.9, 80).astype(np.float32)
    true = np.array([-40.0, 25.0, 10.0, 3.0, -7.0, 120.0])
    vals = orc.ptm_design(lx, ly) @ true
    assert np.allclose(orc.ptm_fit(lx, ly, vals), true, rtol=1e-9, atol=1e-9)


def test_ptm_too_few_samples_raises():
    with pytest.raises(ValueError):
        orc.ptm_fit(np.zeros(5, np.float32), np.zeros(5, np.float32), np.zeros(5, np.int32))


def test_rbf_three_node_hand_case():
    lx = np.array([0.0, 0.6, 0.0], np.float32)
    ly = np.array([0.0, 0.0, 0.8], np.float32)
    a_mat = orc.rbf_matrix(lx, ly)
    assert np.allclose(a_mat, [[0, 0.6, 0.8], [0.6, 0, 1.0], [0.8, 1.0, 0]], atol=1e-7)
    w = orc.rbf_fit(lx, ly, np.array([10, 20, 30]))
    assert np.allclose(a_mat @ w, [10, 20, 30])


def test_rbf_duplicate_nodes_raise():
    lx = np.array([0.1, 0.1, 0.5, -0.3], np.float32)
    ly = np.array([0.2, 0.2, 0.1, 0.4], np.float32)
    with pytest.raises(np.linalg.LinAlgError):
        orc.rbf_fit(lx, ly, np.array([1, 2, 3, 4]))


def test_int_cast_kat():
    g = np.full((1, 1, 1, 1), -172.91)
    assert orc.table_from_grids(g)[0, 0, 0, 0] == -172
    with pytest.raises(ValueError):
        orc.table_from_grids(np.full((1, 1, 1, 1), np.nan))
    with pytest.raises(OverflowError):
        orc.table_from_grids(np.full((1, 1, 1, 1), 3e9))


def test_relight_index_kat(relight_golden):
    g = relight_golden
    for x, l, idx in zip(g["cursor"], g["light"], g["index"]):
        lx, _ = orc.cursor_to_light(int(x), int(x), 400, 400)
        assert lx == l and orc.light_to_index(lx) == idx
    assert g["index"].min() == 0 and g["index"].max() == 99


def test_hsv2bgr_matches_cv2_golden(relight_golden):
    assert np.array_equal(orc.hsv2bgr_u8(relight_golden["hsv"]), relight_golden["bgr"])


def test_hsv2bgr_row_tail_matches_cv2_golden(relight_rows_golden):
    """cvtColor's scalar tail (the last cols % 32 pixels of each row) rounds, the body truncates:
    the reference's default 400-wide ROI has a 16-pixel tail per row."""
    g = relight_rows_golden
    shapes = [k[4:] for k in g if k.startswith("hsv_")]
    assert "12x400" in shapes and len(shapes) >= 5
    tail_hits = 0
    for sh in shapes:
        hsv, bgr = g["hsv_" + sh], g["bgr_" + sh]
        assert np.array_equal(orc.hsv2bgr_u8(hsv), bgr), sh
        tail_hits += np.count_nonzero(orc.hsv2bgr_u8(hsv, simd_lanes=0) != bgr)
    assert tail_hits > 0                      # the goldens do exercise the rounding tail


def test_clamp():
    assert np.array_equal(orc.clamp_values(np.array([-5, 0, 1, 255, 256, 4000])), [0, 0, 1, 255, 255, 255])
