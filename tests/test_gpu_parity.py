"""Parity of the CUDA path (through the C ABI) against the CPU oracle and the committed
reference outputs.  Integer/index work is bit-exact; floating point tolerances are the
ones BASELINE.json's north_star states (PTM coefficients / relit values 1e-4 relative vs
fp64, lx/ly <= 1 ulp of float32) -- the fp64 kernels are in fact held to far tighter
bounds below."""
import numpy as np
import pytest

import rti_oracle as orc

pytestmark = pytest.mark.gpu

PTM_COEF_RTOL = 1e-9        # achieved (fp64 normal equations); north_star allows 1e-4
LIGHT_ULP = 1               # float32 ulps on lx / ly


@pytest.fixture(scope="module")
def torch_cuda(lib):
    import torch
    assert torch.cuda.is_available()
    torch.cuda.set_device(0)
    return torch


def ulp_diff(a, b):
    return np.abs(a.view(np.int32).astype(np.int64) - b.view(np.int32).astype(np.int64))


_KEEPALIVE = []


def dev(torch, a):
    """Upload and keep a reference: callers pass .data_ptr() of the result straight to the
    C ABI, so the tensor must outlive the (asynchronous) kernel that reads it."""
    t = torch.from_numpy(np.ascontiguousarray(a)).cuda()
    _KEEPALIVE.append(t)
    if len(_KEEPALIVE) > 64:
        torch.cuda.synchronize()
        del _KEEPALIVE[:32]
    return t


def gpu_gather(torch, frames, cams, row0=0):
    from rti_b200 import _capi
    n, rows, cols = frames.shape
    d_f, d_c = dev(torch, frames), dev(torch, cams)
    lx = torch.empty((rows * cols, n), dtype=torch.float32, device="cuda")
    ly = torch.empty_like(lx)
    inten = torch.empty((rows * cols, n), dtype=torch.int32, device="cuda")
    _capi.call("rti_gather_soa", d_f.data_ptr(), rows * cols, d_c.data_ptr(), n, row0, rows, cols,
               lx.data_ptr(), ly.data_ptr(), inten.data_ptr(), None)
    torch.cuda.synchronize()
    shape = (rows, cols, n)
    return lx.cpu().numpy().reshape(shape), ly.cpu().numpy().reshape(shape), inten.cpu().numpy().reshape(shape)


def gpu_fit_frames(torch, frames, cams, row0=0):
    from rti_b200 import _capi
    n, rows, cols = frames.shape
    d_f, d_c = dev(torch, frames), dev(torch, cams)
    coef = torch.empty((6, rows * cols), dtype=torch.float64, device="cuda")
    bad = torch.zeros(1, dtype=torch.int32, device="cuda")
    _capi.call("rti_ptm_fit_frames", d_f.data_ptr(), rows * cols, d_c.data_ptr(), n, row0, rows, cols,
               coef.data_ptr(), rows * cols, bad.data_ptr(), None)
    torch.cuda.synchronize()
    return coef, int(bad.item())


def coef_rel_err(coef_gpu, coef_ref):
    """max over pixels of ||a_gpu - a_ref||_inf / ||a_ref||_inf; coef_* are (P, 6)."""
    scale = np.max(np.abs(coef_ref), axis=1, keepdims=True)
    return float(np.max(np.abs(coef_gpu - coef_ref) / scale))


# ------------------------------------------------------------------------------ gather
def test_gather_matches_reference_golden(torch_cuda, ref_small):
    lx, ly, inten = gpu_gather(torch_cuda, ref_small["frames"], ref_small["cams"])
    assert np.array_equal(inten, ref_small["inten"])                  # bit-exact integers
    assert ulp_diff(lx, ref_small["lx"]).max() <= LIGHT_ULP
    assert ulp_diff(ly, ref_small["ly"]).max() <= LIGHT_ULP
    # in practice the fp64 evaluation rounds to the same float32 everywhere
    assert np.count_nonzero(lx != ref_small["lx"]) + np.count_nonzero(ly != ref_small["ly"]) == 0


@pytest.mark.parametrize("rows,cols,n,row0", [(1, 1, 1, 0), (3, 5, 7, 0), (9, 33, 130, 4), (16, 64, 256, 0),
                                              (5, 40, 129, 11), (2, 31, 4, 0),
                                              # n % 8 == 0: the TMA-store kernel (64-frame chunks, 32-pixel CTAs); n % 8 == 4: 16-byte stores
                                              (5, 37, 132, 3), (2, 50, 64, 0), (7, 31, 200, 1), (1, 33, 68, 0),
                                              (3, 100, 500, 2), (4, 45, 136, 1), (2, 70, 504, 0),
                                              # n % 4 != 0 with several 128-frame chunks: coalesced scalar stores
                                              (2, 41, 301, 0), (3, 35, 259, 5)])
def test_gather_ragged_shapes(torch_cuda, rows, cols, n, row0):
    from rti_b200 import synthetic
    frames, cams = synthetic.random_stack(rows, cols, n, seed=rows * 100 + n, rows_total=rows + row0)
    lx, ly, inten = gpu_gather(torch_cuda, frames, cams, row0=row0)
    # oracle evaluates global rows: build the full-height problem and slice
    full = np.zeros((n, rows + row0, cols), np.uint8)
    full[:, row0:] = frames
    o_lx, o_ly, o_in = orc.gather_soa(full, cams)
    assert np.array_equal(inten, o_in[row0:])
    assert ulp_diff(lx, o_lx[row0:]).max() <= LIGHT_ULP
    assert ulp_diff(ly, o_ly[row0:]).max() <= LIGHT_ULP


# ----------------------------------------------------------------------------- PTM fit
def test_ptm_fit_frames_vs_oracle_and_reference(torch_cuda, ref_small):
    coef, bad = gpu_fit_frames(torch_cuda, ref_small["frames"], ref_small["cams"])
    assert bad == 0
    n, rows, cols = ref_small["frames"].shape
    c = coef.cpu().numpy().T                                           # (P, 6)
    ref = orc.ptm_fit_batch(ref_small["lx"], ref_small["ly"], ref_small["inten"]).reshape(-1, 6)
    assert coef_rel_err(c, ref) < PTM_COEF_RTOL
    # against the unmodified reference's grids (its libm powf rounding included): 1e-4 budget
    ax = orc.light_axis()
    for k, (y, x) in enumerate(ref_small["keep"]):
        g = orc.ptm_grid(c[y * cols + x], ax)
        r = ref_small["ptm_grids"][k]
        assert np.max(np.abs(g - r)) <= 1e-6 * np.max(np.abs(r))


@pytest.mark.parametrize("rows,cols,n,row0", [(4, 4, 6, 0), (7, 19, 37, 3), (12, 50, 515, 0), (3, 130, 64, 9)])
def test_ptm_fit_frames_shapes(torch_cuda, rows, cols, n, row0):
    from rti_b200 import synthetic
    frames, cams = synthetic.lambertian_stack(rows, cols, n, seed=n, row0=row0, rows_total=rows + row0)
    coef, bad = gpu_fit_frames(torch_cuda, frames, cams, row0=row0)
    assert bad == 0
    full = np.zeros((n, rows + row0, cols), np.uint8)
    full[:, row0:] = frames
    lx, ly, inten = orc.gather_soa(full, cams)
    ref = orc.ptm_fit_batch(lx[row0:], ly[row0:], inten[row0:]).reshape(-1, 6)
    tol = 1e-7 if n == 6 else PTM_COEF_RTOL                            # N = 6 is exactly determined
    assert coef_rel_err(coef.cpu().numpy().T, ref) < tol


@pytest.mark.parametrize("rows,cols,n", [(6, 21, 203), (5, 27, 64), (3, 50, 33), (9, 17, 130), (2, 129, 97),
                                         (4, 33, 6), (7, 19, 31)])
def test_ptm_fit_soa_matches_frames_path(torch_cuda, rows, cols, n):
    """Every residue of n mod 4: the 16-byte staging is aligned in memory, not in the light index."""
    from rti_b200 import _capi, synthetic
    torch = torch_cuda
    frames, cams = synthetic.lambertian_stack(rows, cols, n, seed=8)
    lx, ly, inten = orc.gather_soa(frames, cams)
    p = rows * cols
    coef = torch.empty((6, p), dtype=torch.float64, device="cuda")
    bad = torch.zeros(1, dtype=torch.int32, device="cuda")
    _capi.call("rti_ptm_fit_soa", dev(torch, lx).data_ptr(), dev(torch, ly).data_ptr(),
               dev(torch, inten).data_ptr(), p, n, coef.data_ptr(), p, bad.data_ptr(), None)
    torch.cuda.synchronize()
    assert int(bad.item()) == 0
    ref = orc.ptm_fit_batch(lx, ly, inten).reshape(-1, 6)
    tol = 1e-7 if n == 6 else PTM_COEF_RTOL                            # N = 6 is exactly determined
    assert coef_rel_err(coef.cpu().numpy().T, ref) < tol
    fused, _ = gpu_fit_frames(torch, frames, cams)
    assert coef_rel_err(coef.cpu().numpy().T, fused.cpu().numpy().T) < (1e-7 if n == 6 else 1e-11)


def test_ptm_analytic_kat(torch_cuda):
    """Intensities from a known biquadratic of the per-pixel light vectors are recovered."""
    from rti_b200 import _capi, synthetic
    torch = torch_cuda
    rng = np.random.default_rng(0)
    n, p = 90, 64
    lx = rng.uniform(-0.9, 0.9, (p, n)).astype(np.float32)
    ly = rng.uniform(-0.9, 0.9, (p, n)).astype(np.float32)
    true = np.array([-40.0, 25.0, 10.0, 3.0, -7.0, 120.0])
    vals = np.rint(orc.ptm_design(lx, ly) @ true).astype(np.int32)     # integer intensities
    coef = torch.empty((6, p), dtype=torch.float64, device="cuda")
    _capi.call("rti_ptm_fit_soa", dev(torch, lx).data_ptr(), dev(torch, ly).data_ptr(),
               dev(torch, vals).data_ptr(), p, n, coef.data_ptr(), p, None, None)
    torch.cuda.synchronize()
    ref = orc.ptm_fit_batch(lx, ly, vals)
    assert coef_rel_err(coef.cpu().numpy().T, ref) < PTM_COEF_RTOL
    assert np.max(np.abs(coef.cpu().numpy().T - true)) < 1.0            # rounding noise only


def test_ptm_singular_pixels_are_flagged(torch_cuda):
    """All light samples identical -> rank-1 design matrix: flagged, coefficients NaN."""
    from rti_b200 import _capi
    torch = torch_cuda
    n, p = 12, 3
    lx = np.full((p, n), 0.25, np.float32)
    ly = np.full((p, n), -0.5, np.float32)
    lx[1] = np.linspace(-0.8, 0.8, n)
    ly[1] = np.linspace(0.7, -0.6, n) ** 2                                # pixel 1 is fine
    vals = np.arange(p * n, dtype=np.int32).reshape(p, n)
    coef = torch.empty((6, p), dtype=torch.float64, device="cuda")
    bad = torch.zeros(1, dtype=torch.int32, device="cuda")
    _capi.call("rti_ptm_fit_soa", dev(torch, lx).data_ptr(), dev(torch, ly).data_ptr(),
               dev(torch, vals).data_ptr(), p, n, coef.data_ptr(), p, bad.data_ptr(), None)
    torch.cuda.synchronize()
    assert int(bad.item()) == 2
    c = coef.cpu().numpy()
    assert np.all(np.isnan(c[:, 0])) and np.all(np.isnan(c[:, 2])) and np.all(np.isfinite(c[:, 1]))


def test_ptm_shared_light_path(torch_cuda):
    """Same light vector for every pixel: coef = pinv(design) . I, one contraction."""
    from rti_b200 import _capi
    torch = torch_cuda
    rng = np.random.default_rng(4)
    n, rows, cols = 77, 5, 23
    lu = rng.uniform(-0.9, 0.9, n).astype(np.float32)
    lv = rng.uniform(-0.9, 0.9, n).astype(np.float32)
    frames = rng.integers(0, 256, (n, rows, cols), dtype=np.uint8)
    design = orc.ptm_design(lu, lv)
    u, s, vt = np.linalg.svd(design, full_matrices=False)
    pinv = (vt.T / s) @ u.T                                            # (6, n), the reference's SVD solve
    p = rows * cols
    coef = torch.empty((6, p), dtype=torch.float64, device="cuda")
    _capi.call("rti_ptm_fit_shared", dev(torch, frames).data_ptr(), p, dev(torch, pinv).data_ptr(), n, p,
               coef.data_ptr(), p, None)
    torch.cuda.synchronize()
    lx = np.broadcast_to(lu, (rows, cols, n))
    ly = np.broadcast_to(lv, (rows, cols, n))
    inten = np.transpose(frames, (1, 2, 0)).astype(np.int32)
    ref = orc.ptm_fit_batch(lx, ly, inten).reshape(-1, 6)
    assert coef_rel_err(coef.cpu().numpy().T, ref) < PTM_COEF_RTOL


# ---------------------------------------------------------------------- grid evaluation
def _random_coef(p, seed=0):
    rng = np.random.default_rng(seed)
    return rng.normal(0, 60, (6, p)) + np.array([0, 0, 0, 0, 0, 120.0])[:, None]


@pytest.mark.parametrize("p,g_step", [(36, 0.02), (1, 0.02), (1027, 0.1), (4096, 0.25), (255, 0.02)])
def test_ptm_table_vs_oracle(torch_cuda, p, g_step):
    from rti_b200 import _capi
    torch = torch_cuda
    coef = _random_coef(p, seed=p)
    ax = orc.light_axis(g_step)
    g = ax.size
    table = torch.empty((g, g, p), dtype=torch.int32, device="cuda")
    bad = torch.zeros(1, dtype=torch.int32, device="cuda")
    _capi.call("rti_ptm_table", dev(torch, coef).data_ptr(), p, p, dev(torch, ax).data_ptr(), g,
               table.data_ptr(), p, bad.data_ptr(), None)
    torch.cuda.synchronize()
    assert int(bad.item()) == 0
    grids = orc.ptm_grid(coef.T, ax)                                    # (p, g, g)
    ref = orc.table_from_grids(grids.reshape(p, 1, g, g))[:, :, :, 0]   # (g, g, p)
    got = table.cpu().numpy()
    diff = got != ref
    # the regrouped fp64 evaluation may flip the truncation only on an integer boundary
    gt = np.transpose(grids, (1, 2, 0))
    assert np.all(np.abs(gt[diff] - np.rint(gt[diff])) < 1e-9)
    assert np.count_nonzero(diff) <= 1e-6 * diff.size + 1
    assert np.all(np.abs(got[diff] - ref[diff]) <= 1) if diff.any() else True


def test_ptm_table_reference_golden(torch_cuda, ref_small):
    """frames -> fused fit -> table equals the unmodified reference's prepare_images_data."""
    from rti_b200 import _capi
    torch = torch_cuda
    coef, _ = gpu_fit_frames(torch, ref_small["frames"], ref_small["cams"])
    ax = orc.light_axis()
    p = coef.shape[1]
    table = torch.empty((100, 100, p), dtype=torch.int32, device="cuda")
    _capi.call("rti_ptm_table", coef.data_ptr(), p, p, dev(torch, ax).data_ptr(), 100, table.data_ptr(), p,
               None, None)
    torch.cuda.synchronize()
    ref = ref_small["table_ptm"].reshape(100, 100, p)
    diff = table.cpu().numpy() != ref
    assert np.count_nonzero(diff) == 0


def test_ptm_table_flags_unrepresentable(torch_cuda):
    from rti_b200 import _capi
    torch = torch_cuda
    coef = _random_coef(8)
    coef[5, 3] = np.nan
    coef[0, 6] = 5e9
    ax = orc.light_axis()
    table = torch.empty((100, 100, 8), dtype=torch.int32, device="cuda")
    bad = torch.zeros(1, dtype=torch.int32, device="cuda")
    _capi.call("rti_ptm_table", dev(torch, coef).data_ptr(), 8, 8, dev(torch, ax).data_ptr(), 100,
               table.data_ptr(), 8, bad.data_ptr(), None)
    torch.cuda.synchronize()
    assert int(bad.item()) >= 1


def test_ptm_grids_vs_oracle(torch_cuda):
    from rti_b200 import _capi
    torch = torch_cuda
    p = 19
    coef = _random_coef(p, seed=3)
    ax = orc.light_axis()
    grids = torch.empty((p, 100, 100), dtype=torch.float64, device="cuda")
    _capi.call("rti_ptm_grids", dev(torch, coef).data_ptr(), p, p, dev(torch, ax).data_ptr(), 100,
               grids.data_ptr(), None)
    torch.cuda.synchronize()
    ref = orc.ptm_grid(coef.T, ax)
    assert np.max(np.abs(grids.cpu().numpy() - ref)) <= 1e-12 * np.max(np.abs(ref))


@pytest.mark.parametrize("g", [1, 7, 255, 256, 257, 300])
def test_ptm_grids_any_axis_length(torch_cuda, g):
    """The kernel walks directions d = v g + u in steps of its CTA size with a (v, u) carry: axis lengths
    below, at and above the CTA size, checked against the oracle's grid evaluation."""
    from rti_b200 import _capi
    torch = torch_cuda
    p = 5
    coef = _random_coef(p, seed=g)
    ax = np.linspace(-1.0, 1.0, g) if g > 1 else np.array([0.25])
    grids = torch.full((p, g, g), np.nan, dtype=torch.float64, device="cuda")
    _capi.call("rti_ptm_grids", dev(torch, coef).data_ptr(), p, p, dev(torch, ax).data_ptr(), g,
               grids.data_ptr(), None)
    torch.cuda.synchronize()
    ref = orc.ptm_grid(coef.T, ax)
    got = grids.cpu().numpy()
    assert got.shape == ref.shape and np.all(np.isfinite(got))
    assert np.max(np.abs(got - ref)) <= 1e-12 * max(np.max(np.abs(ref)), 1.0)


def test_grids_to_table_transposes_and_truncates(torch_cuda):
    from rti_b200 import _capi
    torch = torch_cuda
    rng = np.random.default_rng(1)
    p, g = 45, 10
    grids = rng.normal(0, 300, (p, g, g))
    grids[0, 0, 0] = -172.91
    table = torch.empty((g, g, p), dtype=torch.int32, device="cuda")
    bad = torch.zeros(1, dtype=torch.int32, device="cuda")
    _capi.call("rti_grids_to_table", dev(torch, grids).data_ptr(), p, g, table.data_ptr(), p, bad.data_ptr(), None)
    torch.cuda.synchronize()
    ref = orc.table_from_grids(grids.reshape(p, 1, g, g))[:, :, :, 0]
    assert np.array_equal(table.cpu().numpy(), ref) and table[0, 0, 0].item() == -172
    assert int(bad.item()) == 0
    grids[3, 2, 1] = np.inf
    _capi.call("rti_grids_to_table", dev(torch, grids).data_ptr(), p, g, table.data_ptr(), p, bad.data_ptr(), None)
    torch.cuda.synchronize()
    assert int(bad.item()) >= 1


# --------------------------------------------------------------------------- relighting
def test_relight_reproduces_table_at_grid_nodes(torch_cuda, relight_golden):
    from rti_b200 import _capi, relighting
    torch = torch_cuda
    rows = cols = 64
    p = rows * cols
    coef = _random_coef(p, seed=11)
    ax = orc.light_axis()
    d_coef = dev(torch, coef)
    table = torch.empty((100, 100, p), dtype=torch.int32, device="cuda")
    _capi.call("rti_ptm_table", d_coef.data_ptr(), p, p, dev(torch, ax).data_ptr(), 100, table.data_ptr(), p,
               None, None)
    hsv = relight_golden["hsv"]                                          # (64, 64, 3)
    rl = relighting.Relighter(d_coef, rows, cols, hsv_roi=hsv)
    rl_v = relighting.Relighter(d_coef, rows, cols)
    tab = table.cpu().numpy().reshape(100, 100, rows, cols)
    for (x, y) in [(0, 0), (200, 200), (399, 13), (123, 321), (400, 400)]:
        lx, ly = orc.cursor_to_light(x, y, 400, 400)
        ref_bgr = orc.relight_from_table(tab, hsv, lx, ly)
        got = rl.render_cursor(x, y, shape=(400, 400)).cpu().numpy()
        assert np.array_equal(got, ref_bgr)
        v = rl_v.render_cursor(x, y, shape=(400, 400)).cpu().numpy()
        assert np.array_equal(v, orc.clamp_values(tab[orc.light_to_index(ly)][orc.light_to_index(lx)]))
        assert np.array_equal(relighting.relight_from_table(
            tab[orc.light_to_index(ly)][orc.light_to_index(lx)], hsv), ref_bgr)


def test_hsv2bgr_device_row_tail_matches_cv2_golden(torch_cuda, relight_rows_golden):
    """Widths with a scalar tail (400 is the reference's ROI_DIAMETER): table-lookup and PTM relight
    kernels against cv2.cvtColor's own output (interactive_relighting.py:37-38)."""
    from rti_b200 import relighting
    torch = torch_cuda
    g = relight_rows_golden
    for key in [k for k in g if k.startswith("hsv_")]:
        hsv, bgr = g[key], g["bgr_" + key[4:]]
        rows, cols = hsv.shape[:2]
        out = relighting.relight_from_table(hsv[:, :, 2].astype(np.int32), hsv)
        assert np.array_equal(out, bgr), key
        # constant-polynomial coefficient map whose value is the V channel: same image through relight_kernel
        coef = np.zeros((6, rows * cols))
        coef[5] = hsv[:, :, 2].reshape(-1) + 0.25
        for prec in ("f64", "f32"):
            img = relighting.Relighter(dev(torch, coef), rows, cols, hsv_roi=hsv, precision=prec).render(0.3, -0.6)
            assert np.array_equal(img.cpu().numpy(), bgr), (key, prec)


def test_relight_byte_streams_at_odd_addresses(torch_cuda, relight_rows_golden):
    """The HSV / BGR byte streams are read and written as words when both are 4-byte aligned; any other
    address (a view into a larger buffer) takes the byte path and gives the same image."""
    from rti_b200 import _capi
    torch = torch_cuda
    g = relight_rows_golden
    for key, off in (("hsv_5x399", 1), ("hsv_12x400", 2), ("hsv_9x37", 3)):
        hsv, bgr = g[key], g["bgr_" + key[4:]]
        rows, cols = hsv.shape[:2]
        n = rows * cols
        d_img = dev(torch, hsv[:, :, 2].astype(np.int32).reshape(-1))
        d_hsv = torch.zeros(3 * n + 8, dtype=torch.uint8, device="cuda")
        d_hsv[off:off + 3 * n] = torch.from_numpy(hsv.reshape(-1)).cuda()
        d_out = torch.zeros(3 * n + 8, dtype=torch.uint8, device="cuda")
        _capi.call("rti_relight_from_image", d_img.data_ptr(), n, cols, d_hsv.data_ptr() + off,
                   d_out.data_ptr() + off, None)
        torch.cuda.synchronize()
        out = d_out.cpu().numpy()
        assert np.array_equal(out[off:off + 3 * n].reshape(rows, cols, 3), bgr), key
        assert not out[:off].any() and not out[off + 3 * n:].any()       # nothing written outside


def test_relight_f32_map_within_one(torch_cuda):
    from rti_b200 import relighting
    torch = torch_cuda
    rows = cols = 50
    coef = _random_coef(rows * cols, seed=2)
    exact = relighting.Relighter(dev(torch, coef), rows, cols).render(0.37, -0.81).cpu().numpy().astype(int)
    fast = relighting.Relighter(dev(torch, coef), rows, cols, precision="f32").render(0.37, -0.81).cpu().numpy()
    d = np.abs(exact - fast.astype(int))
    assert d.max() <= 1 and np.count_nonzero(d) <= 0.01 * d.size


def test_hsv2bgr_device_matches_cv2_golden(torch_cuda, relight_golden):
    """Every V value through the device HSV->BGR against the committed cv2 output."""
    from rti_b200 import relighting
    hsv = relight_golden["hsv"]
    out = relighting.relight_from_table(hsv[:, :, 2].astype(np.int32), hsv)
    assert np.array_equal(out, relight_golden["bgr"])


# ---------------------------------------------------------------- call surface, end to end
def test_call_surface_matches_reference_outputs(torch_cuda, ref_small, monkeypatch, tmp_path):
    import rti_b200
    from rti_b200 import synthetic
    frames, cams = ref_small["frames"], ref_small["cams"]
    monkeypatch.setattr(rti_b200.constants, "ROI_DIAMETER", frames.shape[1])
    data = rti_b200.compute_intensities(synthetic.as_records(frames, cams))
    assert len(data) == 3 and data[0].shape == ref_small["lx"].shape
    assert np.array_equal(np.asarray(data[2]), ref_small["inten"])
    assert np.array_equal(data[2][1][2], ref_small["inten"][1][2])       # [y][x] -> (N,)
    assert ulp_diff(np.asarray(data[0]), ref_small["lx"]).max() <= LIGHT_ULP
    res = rti_b200.interpolate_intensities(data, interpolate_PTM=True)
    assert len(res) == frames.shape[1]
    for k, (y, x) in enumerate(ref_small["keep"]):
        g = res[y][x]
        assert g.shape == (100, 100) and g.dtype == np.float64
        r = ref_small["ptm_grids"][k]
        assert np.max(np.abs(g - r)) <= 1e-6 * np.max(np.abs(r))
    images = rti_b200.prepare_images_data(res)
    assert len(images) == 100 and len(images[0]) == 100
    assert images[3][7].dtype == np.int32 and images[3][7].flags.writeable
    assert np.array_equal(np.array(images), ref_small["table_ptm"])
    # plain ndarray inputs (what the reference's own compute_intensities returns) take the SoA kernel
    res2 = rti_b200.interpolate_intensities((ref_small["lx"], ref_small["ly"], ref_small["inten"]),
                                            interpolate_PTM=True)
    assert np.array_equal(np.array(rti_b200.prepare_images_data(res2)), ref_small["table_ptm"])
    # nested-list input to prepare_images_data (the reference's interpolate output type)
    nested = res.to_list()
    assert np.array_equal(np.array(rti_b200.prepare_images_data(nested)), ref_small["table_ptm"])
    # first_only = pixel (0, 0)
    d1 = rti_b200.compute_intensities(synthetic.as_records(frames, cams), first_only=True)
    r1 = rti_b200.interpolate_intensities(d1, interpolate_PTM=True, first_only=True)
    i1 = rti_b200.prepare_images_data(r1, first_only=True)
    assert i1[5][6].shape == (1, 1) and i1[5][6][0, 0] == ref_small["table_ptm"][5, 6, 0, 0]


def test_compute_from_storage_writes_reference_pickle(torch_cuda, ref_small, monkeypatch, tmp_path):
    import pickle
    import rti_b200
    from rti_b200 import synthetic
    frames, cams = ref_small["frames"], ref_small["cams"]
    monkeypatch.setattr(rti_b200.constants, "ROI_DIAMETER", frames.shape[1])
    monkeypatch.chdir(tmp_path)
    (tmp_path / "assets").mkdir()
    rti_b200.write_on_file(synthetic.as_records(frames, cams), "assets/frames_results_synth")
    rti_b200.compute(video_name="synth", from_storage=True, interpolate_PTM=True)
    with open("assets/interpolation_results_synth.pickle", "rb") as f:
        table = pickle.load(f)
    assert isinstance(table, list) and isinstance(table[0], list) and table[0][0].dtype == np.int32
    assert np.array_equal(np.array(table), ref_small["table_ptm"])
    values = table[40][60]
    values[values > 255] = 255                                             # the relighter's in-place clamp
    values[values <= 0] = 0


@pytest.mark.parametrize("transfer,limit,threads,pinned", [
    ("wide", None, None, True),          # int32 tiles straight into the caller's pinned table
    ("narrow", None, None, True),        # int16 over PCIe + host widening threads
    ("narrow", None, "3", False),        # pageable destination, image ranges that do not divide evenly
    ("narrow", "100", "2", True),        # most tiles hold a value beyond the (lowered) limit: int32 re-copy
    ("mixed2", None, "2", True),         # every second tile goes as int32 straight into the table, the rest as int16
    ("mixed", None, None, True),         # int16 unless the staging ring is full at that moment
    ("mixed", None, "1", False),         # pageable destination: never int32 tiles
    (None, None, None, True),            # the library's own choice for this host
])
def test_host_pipeline_tiles_match_single_shot(torch_cuda, monkeypatch, transfer, limit, threads, pinned):
    """rti_ptm_compute_host with many small row tiles == device path on the whole slab, in both transfer
    forms (the table that reaches the host is the same int32 table either way)."""
    from rti_b200 import _capi, synthetic
    torch = torch_cuda
    mix_every = None
    if transfer == "mixed2":
        transfer, mix_every = "mixed", "2"
    for name, val in (("RTI_HOST_TRANSFER", transfer), ("RTI_HOST_NARROW_LIMIT", limit), ("RTI_HOST_THREADS", threads),
                      ("RTI_HOST_MIX_EVERY", mix_every)):
        if val is None:
            monkeypatch.delenv(name, raising=False)
        else:
            monkeypatch.setenv(name, val)
    rows_total, cols, n, row0, rows = 23, 36, 61, 5, 15
    frames, cams = synthetic.lambertian_stack(rows_total, cols, n, seed=21)
    ax = orc.light_axis(0.1)
    g = ax.size
    p = rows * cols
    h_table = torch.empty((g, g, p), dtype=torch.int32)
    h_coef = torch.empty((6, p), dtype=torch.float64).pin_memory()
    if pinned:
        h_table = h_table.pin_memory()
    h_frames = torch.from_numpy(frames).pin_memory()
    lx, ly, inten = orc.gather_soa(frames, cams)
    ref = orc.ptm_fit_batch(lx[row0:row0 + rows], ly[row0:row0 + rows], inten[row0:row0 + rows]).reshape(-1, 6)
    ref_tab = orc.table_from_grids(orc.ptm_grid(ref, ax).reshape(p, 1, g, g))[:, :, :, 0]
    if limit is not None:
        assert np.abs(ref_tab).max() > int(limit)                         # the re-copy path is really taken
    # tiles of 4 rows (144 px: vector narrowing; last tile 3 rows = 108 px: scalar), one tile of 540 px, 8 + 7 rows
    for tile_rows in (3, 0, 5):
        h_table.fill_(-1)
        h_coef.fill_(0)
        _capi.call("rti_ptm_compute_host", h_frames.data_ptr(), cams.ctypes.data, n, rows_total, cols, row0, rows,
                   ax.ctypes.data, g, h_table.data_ptr(), h_coef.data_ptr(), tile_rows)
        assert coef_rel_err(h_coef.numpy().T, ref) < PTM_COEF_RTOL
        assert np.count_nonzero(h_table.numpy() != ref_tab) == 0
        moved = _capi.load().rti_host_d2h_bytes()
        wide_bytes = 4 * g * g * p + 48 * p
        if transfer == "wide":
            assert moved == wide_bytes
        elif transfer == "narrow" and limit is None or transfer == "mixed" and not pinned:
            assert moved == 2 * g * g * p + 48 * p
        elif mix_every is not None and tile_rows == 3:
            assert 2 * g * g * p + 48 * p < moved < wide_bytes                 # some of the four tiles each way
    _capi.call("rti_destroy")


@pytest.mark.parametrize("transfer,limit,threads,pinned", [
    ("wide", None, "2", False), ("narrow", None, "3", False), ("narrow", None, None, True),
    ("narrow", "50", "2", False), (None, None, None, False),
])
def test_table_tile_to_host_every_crossing(torch_cuda, monkeypatch, transfer, limit, threads, pinned):
    """rti_table_tile_to_host: a strided device tile lands in its place inside a larger host table, bit for bit,
    as int16 or int32 over PCIe, through one or many staging chunks, ordered by an event or by the stream."""
    from rti_b200 import _capi
    torch = torch_cuda
    for name, val in (("RTI_HOST_TRANSFER", transfer), ("RTI_HOST_NARROW_LIMIT", limit), ("RTI_HOST_THREADS", threads)):
        if val is None:
            monkeypatch.delenv(name, raising=False)
        else:
            monkeypatch.setenv(name, val)
    rng = np.random.default_rng(5)
    for n_images, tp, stride, host_stride, pix_off, hi in ((7, 37, 41, 100, 13, 300), (400, 108, 108, 540, 432, 300),
                                                          (3, 16, 16, 16, 0, 40000), (70000, 263, 264, 263, 0, 120)):
        tile = rng.integers(-hi, hi, (n_images, stride)).astype(np.int32)
        d_tile = torch.from_numpy(tile).cuda()
        host = torch.full((n_images, host_stride), -7, dtype=torch.int32)
        if pinned:
            host = host.pin_memory()
        ev = torch.cuda.Event()
        ev.record()
        for ready in (ev.cuda_event, None):
            host.fill_(-7)
            _capi.call("rti_table_tile_to_host", d_tile.data_ptr(), stride, n_images, tp, host.data_ptr(), host_stride,
                       pix_off, ready, torch.cuda.current_stream().cuda_stream)
            want = np.full((n_images, host_stride), -7, dtype=np.int32)
            want[:, pix_off:pix_off + tp] = tile[:, :tp]
            assert np.array_equal(host.numpy(), want)
            moved = _capi.load().rti_host_d2h_bytes()
            fits = hi <= 32768 and (limit is None or hi <= int(limit))
            if transfer == "narrow":
                assert moved == n_images * tp * (2 if fits else 4)
            elif transfer == "wide":
                assert moved == n_images * tp * 4
    _capi.call("rti_destroy")


@pytest.mark.parametrize("transfer", ["wide", "narrow", None])
def test_table_tile_to_file_writes_rows_at_their_offsets(torch_cuda, monkeypatch, tmp_path, transfer):
    """rti_table_tile_to_file: rows of every image land at h_offsets[i] + 4 pix_off of the file and nothing else is
    touched; a failing descriptor is an error, not a silent success."""
    import os
    from rti_b200 import _capi
    torch = torch_cuda
    if transfer is None:
        monkeypatch.delenv("RTI_HOST_TRANSFER", raising=False)
    else:
        monkeypatch.setenv("RTI_HOST_TRANSFER", transfer)
    rng = np.random.default_rng(9)
    n_images, rows, cols, gap = 300, 7, 36, 13
    per = rows * cols * 4 + gap
    offsets = (5 + per * np.arange(n_images)).astype(np.int64)
    path = str(tmp_path / "t.bin")
    want = np.full(5 + per * n_images, 0xEE, dtype=np.uint8)
    fd = os.open(path, os.O_RDWR | os.O_CREAT | os.O_TRUNC, 0o644)
    try:
        os.pwrite(fd, want.tobytes(), 0)
        for r0, tr, hi in ((0, 3, 200), (3, 4, 70000)):                   # second band: values beyond int16
            tile = rng.integers(-hi, hi, (n_images, tr * cols + 4)).astype(np.int32)      # strided on the device
            d_tile = torch.from_numpy(tile).cuda()
            _capi.call("rti_table_tile_to_file", d_tile.data_ptr(), tile.shape[1], n_images, tr * cols, fd,
                       offsets.ctypes.data, r0 * cols, None, torch.cuda.current_stream().cuda_stream)
            for i in range(n_images):
                a = offsets[i] + 4 * r0 * cols
                want[a:a + 4 * tr * cols] = tile[i, :tr * cols].view(np.uint8)
        got = np.frombuffer(os.pread(fd, want.size, 0), dtype=np.uint8)
        assert np.array_equal(got, want)
    finally:
        os.close(fd)
    d_tile = torch.zeros((4, 8), dtype=torch.int32, device="cuda")
    with pytest.raises(_capi.RtiError):
        _capi.call("rti_table_tile_to_file", d_tile.data_ptr(), 8, 4, 8, 1 << 20, offsets.ctypes.data, 0, None, None)
    _capi.call("rti_destroy")


def _fit_soa(torch, lx, ly, inten):
    from rti_b200 import _capi
    p, n = lx.shape
    coef = torch.empty((6, p), dtype=torch.float64, device="cuda")
    bad = torch.zeros(1, dtype=torch.int32, device="cuda")
    _capi.call("rti_ptm_fit_soa", dev(torch, lx).data_ptr(), dev(torch, ly).data_ptr(), dev(torch, inten).data_ptr(),
               p, n, coef.data_ptr(), p, bad.data_ptr(), None)
    torch.cuda.synchronize()
    return coef.cpu().numpy().T, int(bad.item())


def test_ptm_fit_narrow_light_cone_and_near_collinear_lights(torch_cuda):
    """the kernels solve the normal equations where the reference takes an SVD of the design
    matrix (analysis.py:294-298).  (a) a narrow light cone (|l| < 0.1: badly SCALED columns) keeps
    north_star's 1e-4 with a wide margin -- Cholesky is insensitive to column scaling; (b) lights on a
    line plus jitter (genuinely ill-conditioned, cond(design) >~ 1e6) are reported as singular instead
    of coming back with half their digits; (c) in between, coefficients hold 1e-4."""
    torch = torch_cuda
    rng = np.random.default_rng(12)
    p, n = 64, 120
    inten = rng.integers(0, 256, (p, n)).astype(np.int32)
    # (a) narrow cone
    lx = rng.uniform(-0.1, 0.1, (p, n)).astype(np.float32)
    ly = rng.uniform(-0.1, 0.1, (p, n)).astype(np.float32)
    got, bad = _fit_soa(torch, lx, ly, inten)
    assert bad == 0
    ref = orc.ptm_fit_batch(lx, ly, inten)
    assert coef_rel_err(got, ref) < 1e-7
    # (b) lights on a line, jitter 1e-7: cond(design) ~ 1e7 or worse
    t = rng.uniform(-0.9, 0.9, (p, n))
    lx = t.astype(np.float32)
    ly = (0.5 * t + 1e-7 * rng.standard_normal((p, n))).astype(np.float32)
    conds = np.array([np.linalg.cond(orc.ptm_design(lx[i], ly[i])) for i in range(p)])
    assert conds.min() > 3e6
    got, bad = _fit_soa(torch, lx, ly, inten)
    assert bad == p and np.all(np.isnan(got))
    # (c) jitter 1e-2: cond(design) ~ 1e4, solvable within the budget
    ly = (0.5 * t + 1e-2 * rng.standard_normal((p, n))).astype(np.float32)
    conds = np.array([np.linalg.cond(orc.ptm_design(lx[i], ly[i])) for i in range(p)])
    assert conds.max() < 1e5
    got, bad = _fit_soa(torch, lx, ly, inten)
    assert bad == 0
    assert coef_rel_err(got, orc.ptm_fit_batch(lx, ly, inten)) < 1e-4


# ------------------------------------------------------- properties at BASELINE sizes
def test_full_size_properties(torch_cuda):
    """Config 3 geometry on a 1024-column slab with N=500: size-independent properties.
    (a) row-slab independence: fitting rows [r0, r1) alone equals the same rows of a larger fit;
    (b) linearity of the fit in the intensities: coef(I1 + I2) = coef(I1) + coef(I2);
    (c) the table image of the grid node (lu, lv) equals the relit frame for that direction."""
    from rti_b200 import _capi, relighting, synthetic
    torch = torch_cuda
    rows, cols, n = 64, 1024, 500
    rng = np.random.default_rng(0)
    cams = synthetic.camera_positions(1024, cols, n, rng)
    f1 = rng.integers(0, 128, (n, rows, cols), dtype=np.uint8)
    f2 = rng.integers(0, 128, (n, rows, cols), dtype=np.uint8)
    c1, b1 = gpu_fit_frames(torch, f1, cams, row0=200)
    c2, _ = gpu_fit_frames(torch, f2, cams, row0=200)
    c12, _ = gpu_fit_frames(torch, f1 + f2, cams, row0=200)
    assert b1 == 0
    scale = c12.abs().max().item()
    assert (c1 + c2 - c12).abs().max().item() < 1e-10 * scale            # (b)
    sub, _ = gpu_fit_frames(torch, f1[:, 16:32], cams, row0=216)
    assert torch.equal(sub, c1.view(6, rows, cols)[:, 16:32].reshape(6, -1))   # (a) bit-identical
    # sampled oracle check at full N
    lx, ly, inten = orc.gather_soa(np.concatenate([np.zeros((n, 200, 8), np.uint8), f1[:, :, :8]], axis=1), cams)
    ref = orc.ptm_fit_batch(lx[200:203], ly[200:203], inten[200:203]).reshape(-1, 6)
    got = c1.view(6, rows, cols)[:, :3, :8].reshape(6, -1).cpu().numpy().T
    assert coef_rel_err(got, ref) < PTM_COEF_RTOL
    ax = orc.light_axis()
    p = rows * cols
    table = torch.empty((100, 100, p), dtype=torch.int32, device="cuda")
    _capi.call("rti_ptm_table", c1.data_ptr(), p, p, dev(torch, ax).data_ptr(), 100, table.data_ptr(), p, None, None)
    rl = relighting.Relighter(c1, rows, cols)
    for v, u in [(0, 0), (99, 99), (50, 17), (3, 88)]:
        frame = rl.render(ax[u], ax[v])
        assert torch.equal(frame.view(-1).to(torch.int32), table[v, u].clamp(0, 255))   # (c)


def test_fused_kernel_whole_slab_against_independent_least_squares(torch_cuda):
    """Every pixel of a 32 x 1024 slab at the bench's light count (N = 500), not a sample: the fused
    gather + fit + table kernel (rti_ptm_fit_table_frames, the headline kernel, two frame chunks per tile)
    against an independent fp64 solution of the same least-squares problems -- the oracle's light vectors
    and design matrix (analysis.py:228-232, 282-290), solved by torch's batched QR on the GPU instead of the
    kernel's normal equations -- and the table against the polynomial of those coefficients
    (analysis.py:303-315, 404-407), every one of the 32768 x 10^4 entries."""
    from rti_b200 import _capi, synthetic
    torch = torch_cuda
    rows, cols, n, row0 = 32, 1024, 500, 700
    rng = np.random.default_rng(31)
    cams = synthetic.camera_positions(1024, cols, n, rng)
    frames = rng.integers(0, 256, (n, rows, cols), dtype=np.uint8)
    p = rows * cols
    ax = orc.light_axis()
    d_f, d_c, d_ax = dev(torch, frames), dev(torch, cams), dev(torch, ax)
    coef = torch.empty((6, p), dtype=torch.float64, device="cuda")
    table = torch.empty((100, 100, p), dtype=torch.int32, device="cuda")
    bad = torch.zeros(2, dtype=torch.int32, device="cuda")
    _capi.call("rti_ptm_fit_table_frames", d_f.data_ptr(), p, d_c.data_ptr(), n, row0, rows, cols, d_ax.data_ptr(), 100,
               coef.data_ptr(), p, table.data_ptr(), p, bad.data_ptr(), None)
    torch.cuda.synchronize()
    assert bad.sum().item() == 0
    worst_coef, off = 0.0, 0
    gu = torch.from_numpy(ax).cuda()
    lu = gu[None, None, :]
    lv = gu[None, :, None]
    for y in range(rows):
        lx, ly, inten = orc.gather_soa(frames[:, y:y + 1], cams, row0=row0 + y)      # (1, cols, n)
        design = torch.from_numpy(np.stack([orc.ptm_design(lx[0, x], ly[0, x]) for x in range(0, cols, 1)])).cuda()
        rhs = torch.from_numpy(inten[0].astype(np.float64)).cuda()[:, :, None]
        sol = torch.linalg.lstsq(design, rhs, driver="gels").solution[:, :, 0]       # (cols, 6)
        got = coef[:, y * cols:(y + 1) * cols].T
        worst_coef = max(worst_coef, ((got - sol).abs().amax(dim=1) / sol.abs().amax(dim=1)).max().item())
        a = got                                                                       # the kernel's own coefficients
        val = (a[:, 0, None, None] * lu * lu + a[:, 1, None, None] * lv * lv + a[:, 2, None, None] * lu * lv
               + a[:, 3, None, None] * lu + a[:, 4, None, None] * lv + a[:, 5, None, None])   # (cols, 100 v, 100 u)
        t = table[:, :, y * cols:(y + 1) * cols].permute(2, 0, 1)
        diff = t != torch.trunc(val).to(torch.int32)
        # a differently grouped fp64 sum may cross an integer only when the value sits on it
        off += int((diff & ((val - torch.round(val)).abs() > 1e-9)).sum().item())
    assert worst_coef < PTM_COEF_RTOL, worst_coef
    assert off == 0


# ------------------------------------------------------------------------------ linear RBF
RBF_GRID_RTOL = 1e-8        # achieved (fp64 Cholesky of the reduced SPD system); north_star allows 1e-3


def gpu_rbf_fit(torch, lx, ly, inten):
    """lx, ly (P, n) float32, inten (P, n) int32 -> weights (P, n) float64, bad count."""
    from rti_b200 import _capi
    lib = _capi.load()
    p, n = lx.shape
    d_lx, d_ly, d_in = dev(torch, lx), dev(torch, ly), dev(torch, inten)
    w = torch.empty((p, n), dtype=torch.float64, device="cuda")
    bad = torch.zeros(1, dtype=torch.int32, device="cuda")
    batch = min(p, 32)
    nbytes = lib.rti_rbf_workspace_bytes(n, batch)
    work = torch.empty(nbytes, dtype=torch.uint8, device="cuda")
    _capi.call("rti_rbf_fit_soa", d_lx.data_ptr(), d_ly.data_ptr(), d_in.data_ptr(), p, n, w.data_ptr(),
               work.data_ptr(), nbytes, batch, bad.data_ptr(), None)
    torch.cuda.synchronize()
    return w, int(bad.item()), (d_lx, d_ly)


def random_nodes(p, n, seed):
    rng = np.random.default_rng(seed)
    r = np.sqrt(rng.uniform(0, 0.95, (p, n)))
    th = rng.uniform(0, 2 * np.pi, (p, n))
    lx = (r * np.cos(th)).astype(np.float32)
    ly = (r * np.sin(th)).astype(np.float32)
    inten = rng.integers(0, 256, (p, n)).astype(np.int32)
    return lx, ly, inten


@pytest.mark.parametrize("p,n", [(5, 2), (5, 3), (40, 40), (7, 64), (7, 65), (7, 66), (6, 96), (9, 129),
                                 (3, 191), (4, 300), (2, 515), (4, 1000)])
def test_rbf_fit_vs_oracle(torch_cuda, p, n):
    lx, ly, inten = random_nodes(p, n, seed=n)
    w, bad, _ = gpu_rbf_fit(torch_cuda, lx, ly, inten)
    assert bad == 0
    w = w.cpu().numpy()
    # n = 1000 is config 4's light count: cond(A) ~ 1e6-1e7 there, both solvers carry ~cond * 1e-16
    wtol, rtol = (1e-8, 1e-7) if n <= 600 else (1e-7, 1e-6)
    for i in range(p):
        ref = orc.rbf_fit(lx[i], ly[i], inten[i])
        assert np.max(np.abs(w[i] - ref)) <= wtol * np.max(np.abs(ref))
        # the interpolant reproduces the node values (scipy KAT)
        assert np.allclose(orc.rbf_matrix(lx[i], ly[i]) @ w[i], inten[i], atol=rtol)


def test_rbf_fit_and_eval_n1000_vs_oracle(torch_cuda):
    """Config 4's light count through the per-pixel fp64 path: fit + grid evaluation + int32 table
    of 4 pixels with 1000 nodes each against the oracle (analysis.py:249-260, :400-409)."""
    from rti_b200 import _capi
    torch = torch_cuda
    p, n = 4, 1000
    lx, ly, inten = random_nodes(p, n, seed=4096)
    w, bad, (d_lx, d_ly) = gpu_rbf_fit(torch, lx, ly, inten)
    assert bad == 0
    ax = orc.light_axis()
    d_ax = dev(torch, ax)
    grids = torch.empty((p, 100, 100), dtype=torch.float64, device="cuda")
    _capi.call("rti_rbf_grids", d_lx.data_ptr(), d_ly.data_ptr(), w.data_ptr(), p, n, d_ax.data_ptr(), 100,
               grids.data_ptr(), None)
    table = torch.empty((100, 100, p), dtype=torch.int32, device="cuda")
    tb = torch.zeros(1, dtype=torch.int32, device="cuda")
    _capi.call("rti_rbf_table", d_lx.data_ptr(), d_ly.data_ptr(), w.data_ptr(), p, n, d_ax.data_ptr(), 100,
               table.data_ptr(), p, tb.data_ptr(), None)
    torch.cuda.synchronize()
    assert int(tb.item()) == 0
    got = grids.cpu().numpy()
    tab = table.cpu().numpy()
    for i in range(p):
        ref = orc.rbf_grid(lx[i], ly[i], orc.rbf_fit(lx[i], ly[i], inten[i]), ax)
        scale = np.max(np.abs(ref))
        assert np.max(np.abs(got[i] - ref)) <= 1e-7 * scale
        diff = tab[:, :, i] != np.trunc(ref).astype(np.int32)
        assert np.all(np.abs(ref[diff] - np.rint(ref[diff])) < 1e-6 * scale)
        assert np.count_nonzero(diff) <= 2


RBF_FAST_RTOL = 5e-4       # float32 distances, fp64 weights and sums: ~5e-5 typical, 2e-4 worst on white-noise
                           # intensities at N = 500; north_star allows 1e-3 of the grid value


@pytest.mark.parametrize("p,n", [(13, 40), (8, 300), (19, 500), (4, 1000)])
def test_rbf_table_fast_vs_oracle(torch_cuda, p, n):
    """rti_rbf_table_fast (float32 node distances) against scipy's fp64 evaluation (analysis.py:259-260,
    :400-409): every table entry within RBF_FAST_RTOL * max|grid| + 1 of trunc(oracle), and equal to it
    wherever the fp64 value is farther than that from an integer."""
    from rti_b200 import _capi
    torch = torch_cuda
    lx, ly, inten = random_nodes(p, n, seed=77 + n)
    w, bad, (d_lx, d_ly) = gpu_rbf_fit(torch, lx, ly, inten)
    assert bad == 0
    ax = orc.light_axis()
    d_ax = dev(torch, ax)
    stride = p + 3                                        # padded image stride, odd pixel count
    table = torch.full((100, 100, stride), -7, dtype=torch.int32, device="cuda")
    tb = torch.zeros(1, dtype=torch.int32, device="cuda")
    _capi.call("rti_rbf_table_fast", d_lx.data_ptr(), d_ly.data_ptr(), w.data_ptr(), p, n, d_ax.data_ptr(), 100,
               table.data_ptr(), stride, tb.data_ptr(), None)
    torch.cuda.synchronize()
    assert int(tb.item()) == 0
    tab = table.cpu().numpy()
    assert np.all(tab[:, :, p:] == -7)                    # nothing written past the pixels
    for i in range(p):
        ref = orc.rbf_grid(lx[i], ly[i], orc.rbf_fit(lx[i], ly[i], inten[i]), ax)
        tol = RBF_FAST_RTOL * np.max(np.abs(ref))
        exact = np.trunc(ref).astype(np.int32)
        diff = tab[:, :, i] != exact
        assert np.all(np.abs(tab[:, :, i] - exact) <= 1)
        assert np.all(np.abs(ref[diff] - np.rint(ref[diff])) <= tol)


def test_rbf_reference_golden_grids_and_table(torch_cuda, ref_small, ref_rbf_n96):
    from rti_b200 import _capi
    torch = torch_cuda
    ax = orc.light_axis()
    d_ax = dev(torch, ax)
    # N = 96: two pixels against the unmodified reference's _interpolate_RBF output
    g = ref_rbf_n96
    lx = np.stack([g["lx"][y, x] for y, x in g["pix"]])
    ly = np.stack([g["ly"][y, x] for y, x in g["pix"]])
    inten = np.stack([g["inten"][y, x] for y, x in g["pix"]])
    w, bad, (d_lx, d_ly) = gpu_rbf_fit(torch, lx, ly, inten)
    assert bad == 0
    grids = torch.empty((2, 100, 100), dtype=torch.float64, device="cuda")
    _capi.call("rti_rbf_grids", d_lx.data_ptr(), d_ly.data_ptr(), w.data_ptr(), 2, 96, d_ax.data_ptr(), 100,
               grids.data_ptr(), None)
    torch.cuda.synchronize()
    for k in range(2):
        r = g["rbf_grids"][k]
        assert np.max(np.abs(grids[k].cpu().numpy() - r)) <= RBF_GRID_RTOL * np.max(np.abs(r))
    # R = 6, N = 40: the whole int32 table against the reference's prepare_images_data
    s = ref_small
    p = s["lx"].shape[0] * s["lx"].shape[1]
    n = s["lx"].shape[2]
    w, bad, (d_lx, d_ly) = gpu_rbf_fit(torch, s["lx"].reshape(p, n), s["ly"].reshape(p, n), s["inten"].reshape(p, n))
    assert bad == 0
    table = torch.empty((100, 100, p), dtype=torch.int32, device="cuda")
    tb = torch.zeros(1, dtype=torch.int32, device="cuda")
    _capi.call("rti_rbf_table", d_lx.data_ptr(), d_ly.data_ptr(), w.data_ptr(), p, n, d_ax.data_ptr(), 100,
               table.data_ptr(), p, tb.data_ptr(), None)
    torch.cuda.synchronize()
    assert int(tb.item()) == 0
    ref = s["table_rbf"].reshape(100, 100, p)
    diff = table.cpu().numpy() != ref
    # flips only where the reference's fp64 value sits on an integer boundary
    grids = orc.interpolate(s["lx"], s["ly"], s["inten"], ptm=False)
    gt = np.transpose(grids.reshape(p, 100, 100), (1, 2, 0))
    assert np.all(np.abs(gt[diff] - np.rint(gt[diff])) < 1e-6)
    assert np.count_nonzero(diff) <= 1e-5 * diff.size + 1


def test_rbf_duplicate_nodes_are_flagged(torch_cuda):
    """scipy raises LinAlgError (singular matrix) for duplicate nodes; the kernel counts them."""
    lx, ly, inten = random_nodes(4, 50, seed=1)
    lx[1, 17], ly[1, 17] = lx[1, 3], ly[1, 3]           # duplicate pair away from node 0
    lx[2, 9], ly[2, 9] = lx[2, 0], ly[2, 0]             # duplicate of node 0
    w, bad, _ = gpu_rbf_fit(torch_cuda, lx, ly, inten)
    assert bad == 2
    w = w.cpu().numpy()
    assert np.all(np.isfinite(w[0])) and np.all(np.isfinite(w[3]))
    # (the CPU solvers see the same singularity only up to round-off: LAPACK raises or returns
    # ~1e16-sized weights depending on the host, so the oracle is not asserted on here)


@pytest.mark.parametrize("precision", ["fp64", "fast"])
def test_rbf_call_surface(torch_cuda, ref_small, monkeypatch, precision):
    """interpolate_intensities(data) + prepare_images_data in linear-RBF mode against the reference's own
    outputs, on both numerical paths: "fp64" (direct Cholesky, fp64 evaluation) and the default "fast"
    (tile-preconditioned CG with direct refits + float32-distance table; north_star: 1e-3)."""
    import rti_b200
    from rti_b200 import analysis, synthetic
    frames, cams = ref_small["frames"], ref_small["cams"]
    monkeypatch.setattr(rti_b200.constants, "ROI_DIAMETER", frames.shape[1])
    monkeypatch.setattr(analysis, "RBF_PRECISION", precision)
    data = rti_b200.compute_intensities(synthetic.as_records(frames, cams))
    res = rti_b200.interpolate_intensities(data)                       # interpolate_PTM=False: RBF
    assert res.precision == precision and res.path.startswith("rbf:")
    rtol = RBF_GRID_RTOL if precision == "fp64" else 2e-4
    for k, (y, x) in enumerate(ref_small["keep"]):
        r = ref_small["rbf_grids"][k]
        assert np.max(np.abs(res[y][x] - r)) <= rtol * np.max(np.abs(r))
    table = np.array(rti_b200.prepare_images_data(res))
    diff = table != ref_small["table_rbf"]
    if precision == "fp64":
        assert np.count_nonzero(diff) <= 1e-5 * diff.size + 1
    else:
        # entries may flip by one where the fp64 value is within the path's tolerance of an integer
        grids = orc.interpolate(ref_small["lx"], ref_small["ly"], ref_small["inten"], ptm=False)
        gt = np.transpose(grids, (2, 3, 0, 1))
        assert np.all(np.abs(table - ref_small["table_rbf"]) <= 1)
        assert np.all(np.abs(gt[diff] - np.rint(gt[diff])) <= RBF_FAST_RTOL * np.max(np.abs(gt)))
    # singular pixels surface as an ArithmeticError subclass
    lx, ly, inten = (np.array(ref_small[k]) for k in ("lx", "ly", "inten"))
    lx[0, 0, 5], ly[0, 0, 5] = lx[0, 0, 6], ly[0, 0, 6]
    with pytest.raises(rti_b200.RtiSingularError):
        rti_b200.interpolate_intensities((lx, ly, inten))


@pytest.mark.parametrize("precision", ["fast", "fp64"])
def test_rbf_banded_route_equals_whole_roi_route(torch_cuda, monkeypatch, precision):
    """Large ROIs in linear-RBF mode never materialise the (R, R, N) arrays or the weights: gather, fit and
    evaluation run per band of ROI rows while the table is produced (analysis.RBF_BANDED_FRACTION).  Forced on
    here for a 40 x 40 ROI with 3 bands and small table tiles: same table and same res[y][x] as the eager route."""
    import rti_b200
    from rti_b200 import analysis, synthetic
    r, n = 40, 30
    frames, cams = synthetic.lambertian_stack(r, r, n, seed=17, rows_total=1024, row0=100)
    monkeypatch.setattr(rti_b200.constants, "ROI_DIAMETER", r)
    monkeypatch.setattr(analysis, "RBF_PRECISION", precision)
    monkeypatch.setattr(analysis, "TABLE_TILE_BYTES", 16 * r * 100 * 100 * 4)          # 16 ROI rows per table tile
    recs = synthetic.as_records(frames, cams)
    eager = rti_b200.interpolate_intensities(rti_b200.compute_intensities(recs))
    assert not eager.banded
    t_eager = np.array(rti_b200.prepare_images_data(eager))
    monkeypatch.setattr(analysis, "RBF_BANDED_FRACTION", 0.0)
    banded = rti_b200.interpolate_intensities(rti_b200.compute_intensities(recs))
    assert banded.banded and banded.path.startswith("rbf: banded")
    t_banded = np.array(rti_b200.prepare_images_data(banded))
    assert np.array_equal(t_eager, t_banded)
    for y, x in [(0, 0), (15, 39), (16, 0), (39, 39), (33, 7)]:
        assert np.array_equal(eager[y][x], banded[y][x])
    g = banded.grids(15 * r + 30, 25)                                  # a pixel range that crosses a band boundary
    assert np.array_equal(g, eager.grids(15 * r + 30, 25))


@pytest.mark.parametrize("ptm", [True, False])
def test_compute_tiles_table_under_device_memory_cap(torch_cuda, monkeypatch, tmp_path, ptm):
    """compute() evaluates the table in row tiles through a bounded device buffer and
    pwrites them into the reference's pickle layout, so a table larger than the device memory it is allowed
    still comes out right.  ROI 96 x 96 (368 MB table) under a 256 MB torch allocator cap with 24 MB tiles;
    sampled pixels against the oracle, every image against the untiled device table."""
    import pickle
    import rti_b200
    from rti_b200 import analysis, synthetic
    torch = torch_cuda
    r, n = 96, 12 if ptm else 24
    frames, cams = synthetic.lambertian_stack(r, r, n, seed=21, rows_total=1024, row0=300)
    path = str(tmp_path / "frames_results_cap")
    rti_b200.write_on_file(synthetic.as_records(frames, cams), path)
    monkeypatch.setattr(rti_b200.constants, "ROI_DIAMETER", r)
    monkeypatch.setattr(analysis, "TABLE_TILE_BYTES", 24 << 20)
    monkeypatch.chdir(tmp_path)
    (tmp_path / "assets").mkdir()
    total = torch.cuda.get_device_properties(0).total_memory
    # the cap counts what the caching allocator has RESERVED: drop what earlier tests left (tensors kept alive for
    # asynchronous calls can pin a large segment) and allow 256 MB beyond whatever remains
    import gc
    torch.cuda.synchronize()
    del _KEEPALIVE[:]
    gc.collect()
    torch.cuda.empty_cache()
    torch.cuda.set_per_process_memory_fraction(min(1.0, (torch.cuda.memory_reserved() + (256 << 20)) / total))
    try:
        rti_b200.compute("cap", from_storage=True, storage_filepath=path, interpolate_PTM=ptm)
    finally:
        torch.cuda.set_per_process_memory_fraction(1.0)
    with open(tmp_path / "assets" / "interpolation_results_cap.pickle", "rb") as f:
        images = pickle.load(f)
    assert len(images) == 100 and len(images[0]) == 100 and images[99][99].shape == (r, r)
    assert images[3][7].dtype == np.int32 and images[3][7].flags.writeable
    table = np.array(images)                                          # (100, 100, r, r)
    lx, ly, inten = orc.gather_soa(frames, cams)                      # compute() numbers the ROI rows from 0
    ax = orc.light_axis()
    rng = np.random.default_rng(2)
    for _ in range(10):
        y, x = int(rng.integers(r)), int(rng.integers(r))
        if ptm:
            ref = orc.ptm_grid(orc.ptm_fit(lx[y, x], ly[y, x], inten[y, x]), ax)
            assert np.array_equal(table[:, :, y, x], np.trunc(ref).astype(np.int32))
        else:
            ref = orc.rbf_grid(lx[y, x], ly[y, x], orc.rbf_fit(lx[y, x], ly[y, x], inten[y, x]), ax)
            d = table[:, :, y, x] != np.trunc(ref).astype(np.int32)
            assert np.all(np.abs(table[:, :, y, x] - np.trunc(ref)) <= 1)
            assert np.all(np.abs(ref[d] - np.rint(ref[d])) <= RBF_FAST_RTOL * np.max(np.abs(ref)))
    # and the tiled file equals the table evaluated in one piece
    data = rti_b200.compute_intensities(synthetic.as_records(frames, cams))
    res = rti_b200.interpolate_intensities(data, interpolate_PTM=ptm)
    whole = analysis.table_device(res).view(100, 100, r, r).cpu().numpy()
    assert np.array_equal(whole, table)


# ------------------------------------------------------------------------- fused kernel
@pytest.mark.parametrize("rows,cols,n,row0,g_step", [(6, 6, 40, 0, 0.02), (5, 103, 77, 2, 0.1), (3, 257, 31, 0, 0.25),
                                                      (16, 64, 130, 7, 0.02)])
def test_fused_kernel_equals_two_kernel_path(torch_cuda, rows, cols, n, row0, g_step):
    """rti_ptm_fit_table_frames is bit-identical to rti_ptm_fit_frames + rti_ptm_table."""
    from rti_b200 import _capi, synthetic
    torch = torch_cuda
    frames, cams = synthetic.lambertian_stack(rows, cols, n, seed=3 * n, row0=row0, rows_total=rows + row0)
    ax = orc.light_axis(g_step)
    g = ax.size
    p = rows * cols
    d_f, d_c, d_ax = dev(torch, frames), dev(torch, cams), dev(torch, ax)
    coef, bad = gpu_fit_frames(torch, frames, cams, row0=row0)
    table = torch.empty((g, g, p), dtype=torch.int32, device="cuda")
    _capi.call("rti_ptm_table", coef.data_ptr(), p, p, d_ax.data_ptr(), g, table.data_ptr(), p, None, None)
    coef2 = torch.empty_like(coef)
    table2 = torch.full_like(table, -7)
    bad2 = torch.zeros(2, dtype=torch.int32, device="cuda")
    _capi.call("rti_ptm_fit_table_frames", d_f.data_ptr(), p, d_c.data_ptr(), n, row0, rows, cols, d_ax.data_ptr(), g,
               coef2.data_ptr(), p, table2.data_ptr(), p, bad2.data_ptr(), None)
    torch.cuda.synchronize()
    assert bad == 0 and bad2.sum().item() == 0
    assert torch.equal(coef, coef2)
    assert torch.equal(table, table2)


@pytest.mark.parametrize("rows,cols,n,row0,g_step", [
    (5, 103, 77, 2, 0.1),        # 3 tiles, the last with 3 pixels; one frame chunk (chunks hold 256 frames)
    (1, 16, 6, 0, 0.5),          # one partial tile, fewest lights
    (4, 64, 64, 0, 0.25),        # exactly one tile
    (3, 257, 200, 5, 0.25),      # one chunk per tile, partial last tile
    (451, 256, 70, 3, 0.25),     # 451 tiles on <= 148 persistent CTAs: >= 3 tiles each (buffer hand-back)
    (37, 1000, 130, 0, 0.5),     # 145 tiles: fewer tiles than SMs
    (3, 300, 257, 1, 0.5),       # 2 chunks per tile (256 + 1): the ring's second stage, partial tile
    (2, 520, 600, 0, 0.5),       # 3 chunks per tile: the two-stage ring wraps inside a tile
])
def test_fused_ws_kernel_padded_stride(torch_cuda, rows, cols, n, row0, g_step):
    """The warp-specialised persistent kernel (taken for 16-byte aligned frame rows) against the
    two-kernel path, bit for bit, with a padded frame stride so partial tiles are exercised."""
    from rti_b200 import _capi, synthetic
    torch = torch_cuda
    frames, cams = synthetic.lambertian_stack(rows, cols, n, seed=5 * n + rows, row0=row0, rows_total=rows + row0)
    ax = orc.light_axis(g_step)
    g = ax.size
    p = rows * cols
    stride = (p + 15) // 16 * 16 + 16
    padded = np.full((n, stride), 255, dtype=np.uint8)
    padded[:, :p] = frames.reshape(n, p)
    d_f, d_c, d_ax = dev(torch, padded), dev(torch, cams), dev(torch, ax)
    coef, bad = gpu_fit_frames(torch, frames, cams, row0=row0)
    table = torch.empty((g, g, p), dtype=torch.int32, device="cuda")
    _capi.call("rti_ptm_table", coef.data_ptr(), p, p, d_ax.data_ptr(), g, table.data_ptr(), p, None, None)
    coef2 = torch.empty_like(coef)
    table2 = torch.full_like(table, -7)
    bad2 = torch.zeros(2, dtype=torch.int32, device="cuda")
    _capi.call("rti_ptm_fit_table_frames", d_f.data_ptr(), stride, d_c.data_ptr(), n, row0, rows, cols, d_ax.data_ptr(),
               g, coef2.data_ptr(), p, table2.data_ptr(), p, bad2.data_ptr(), None)
    torch.cuda.synchronize()
    assert bad == 0 and bad2.sum().item() == 0
    assert torch.equal(coef, coef2)
    assert torch.equal(table, table2)


def oracle_pixels(frames, cams, pix, cols, row0, axis):
    """Oracle coefficients (len(pix), 6) and truncated table columns (G*G, len(pix)) of the flat
    pixel indices `pix` of a (rows, cols) slab whose first row is global row row0."""
    n = frames.shape[0]
    flat = frames.reshape(n, -1)
    coef = np.empty((len(pix), 6))
    for i, q in enumerate(pix):
        y, x = divmod(int(q), cols)
        lx, ly, inten = orc.gather_soa(flat[:, q].reshape(n, 1, 1), cams, row0=row0 + y, col0=x)
        coef[i] = orc.ptm_fit_batch(lx, ly, inten).reshape(6)
    tab = orc.table_from_grids(orc.ptm_grid(coef, axis).reshape(len(pix), 1, axis.size, axis.size))
    return coef, tab.reshape(axis.size * axis.size, len(pix))


@pytest.mark.parametrize("n", [257, 500, 512, 513, 1000])
def test_fused_ws_kernel_multi_chunk_vs_oracle(torch_cuda, n):
    """The headline kernel in the regime it is benchmarked in: more lights than one 256-frame
    chunk of its cp.async ring (n = 500 is the bench's 256 + 244, n = 1000 config 4's four chunks),
    586 tiles on <= 148 persistent CTAs (>= 3 tiles each, so the producer cursor crosses tile
    boundaries mid-ring) and a partial last tile.  Checked against the CPU oracle on sampled pixels
    (tile edges included) and, for every pixel, bit for bit against the two-kernel path."""
    from rti_b200 import _capi, synthetic
    torch = torch_cuda
    rows, cols, row0 = 600, 250, 11
    p = rows * cols                                 # 150 000 = 585 full tiles + 240 pixels, 16-byte aligned rows
    frames, cams = synthetic.random_stack(rows, cols, n, seed=n, rows_total=rows + row0)
    ax = orc.light_axis(0.25)
    g = ax.size
    d_f, d_c, d_ax = dev(torch, frames), dev(torch, cams), dev(torch, ax)
    coef = torch.empty((6, p), dtype=torch.float64, device="cuda")
    table = torch.full((g, g, p), -7, dtype=torch.int32, device="cuda")
    bad = torch.zeros(2, dtype=torch.int32, device="cuda")
    _capi.call("rti_ptm_fit_table_frames", d_f.data_ptr(), p, d_c.data_ptr(), n, row0, rows, cols, d_ax.data_ptr(), g,
               coef.data_ptr(), p, table.data_ptr(), p, bad.data_ptr(), None)
    torch.cuda.synchronize()
    assert bad.sum().item() == 0
    rng = np.random.default_rng(n)
    pix = np.unique(np.concatenate([[0, 255, 256, 511, 512, 585 * 256 - 1, 585 * 256, p - 1],
                                    rng.integers(0, p, 56)]))
    ref_coef, ref_tab = oracle_pixels(frames, cams, pix, cols, row0, ax)
    got = coef[:, torch.from_numpy(pix).cuda()].cpu().numpy().T
    assert coef_rel_err(got, ref_coef) < PTM_COEF_RTOL
    got_tab = table.view(g * g, p)[:, torch.from_numpy(pix).cuda()].cpu().numpy()
    assert np.array_equal(got_tab, ref_tab)
    # every pixel: the two-kernel path (itself held to the oracle above and at n = 500 below)
    coef2, bad2 = gpu_fit_frames(torch, frames, cams, row0=row0)
    table2 = torch.empty_like(table)
    _capi.call("rti_ptm_table", coef2.data_ptr(), p, p, d_ax.data_ptr(), g, table2.data_ptr(), p, None, None)
    torch.cuda.synchronize()
    assert bad2 == 0
    assert torch.equal(coef, coef2)
    assert torch.equal(table, table2)


def test_fused_kernel_reference_golden(torch_cuda, ref_small):
    from rti_b200 import _capi
    torch = torch_cuda
    frames, cams = ref_small["frames"], ref_small["cams"]
    n, rows, cols = frames.shape
    p = rows * cols
    d_f, d_c, d_ax = dev(torch, frames), dev(torch, cams), dev(torch, orc.light_axis())
    table = torch.empty((100, 100, p), dtype=torch.int32, device="cuda")
    bad = torch.zeros(2, dtype=torch.int32, device="cuda")
    _capi.call("rti_ptm_fit_table_frames", d_f.data_ptr(), p, d_c.data_ptr(), n, 0, rows, cols, d_ax.data_ptr(), 100,
               None, 0, table.data_ptr(), p, bad.data_ptr(), None)
    torch.cuda.synchronize()
    assert bad.sum().item() == 0
    assert np.array_equal(table.cpu().numpy().reshape(100, 100, rows, cols), ref_small["table_ptm"])


def test_fused_ws_kernel_stays_inside_its_buffers(torch_cuda):
    """compute-sanitizer is not available on the GPU pool: guard bands around every output of the
    persistent kernel (partial last tile, padded strides) must come back untouched, and the frame
    reads must not need anything past the last valid byte (the stack is allocated exactly)."""
    from rti_b200 import _capi, synthetic
    torch = torch_cuda
    rows, cols, n, g = 7, 131, 70, 8              # 917 pixels: 4 tiles, last one with 149 pixels
    p = rows * cols
    frames, cams = synthetic.lambertian_stack(rows, cols, n, seed=8)
    stride = (p + 15) // 16 * 16
    flat = torch.zeros(n * stride - (stride - p), dtype=torch.uint8, device="cuda")   # ends at the last valid byte
    for i in range(n):
        flat[i * stride:i * stride + p] = torch.from_numpy(frames[i].reshape(-1)).cuda()
    ax = orc.light_axis(0.25)
    d_c, d_ax = dev(torch, cams), dev(torch, ax)
    guard = 64
    plane, image = p + 5, p + 12                  # strides larger than the pixel count
    coef = torch.full((guard + 6 * plane + guard,), -777.0, dtype=torch.float64, device="cuda")
    table = torch.full((guard + g * g * image + guard,), -777, dtype=torch.int32, device="cuda")
    bad = torch.zeros(2 + 2 * guard, dtype=torch.int32, device="cuda")
    _capi.call("rti_ptm_fit_table_frames", flat.data_ptr(), stride, d_c.data_ptr(), n, 0, rows, cols, d_ax.data_ptr(),
               g, coef.data_ptr() + 8 * guard, plane, table.data_ptr() + 4 * guard, image,
               bad.data_ptr() + 4 * guard, None)
    torch.cuda.synchronize()
    assert bad.sum().item() == 0
    c = coef.cpu().numpy()
    t = table.cpu().numpy()
    assert np.all(c[:guard] == -777.0) and np.all(c[-guard:] == -777.0)
    assert np.all(t[:guard] == -777) and np.all(t[-guard:] == -777)
    cv = c[guard:-guard].reshape(6, plane)
    tv = t[guard:-guard].reshape(g * g, image)
    assert np.all(cv[:, p:] == -777.0) and np.all(tv[:, p:] == -777)        # stride padding untouched
    lx, ly, inten = orc.gather_soa(frames, cams)
    ref = orc.ptm_fit_batch(lx, ly, inten).reshape(p, 6).T
    assert np.abs(cv[:, :p] - ref).max() <= PTM_COEF_RTOL * np.abs(ref).max()
    want = orc.table_from_grids(orc.ptm_grid(ref.T.reshape(rows, cols, 6), ax)).reshape(g * g, p)
    assert np.array_equal(tv[:, :p], want)


# ------------------------------------------------------------------ per-pixel validity masks
def test_ptm_masked_fit_equals_reference_on_masked_subarrays(torch_cuda):
    """Extension named by the north star (per-pixel validity masks): the oracle is the reference's
    per-pixel fit called with only the samples whose intensity lies in [lo, hi]."""
    from rti_b200 import _capi, synthetic
    torch = torch_cuda
    rows, cols, n, lo, hi = 9, 21, 60, 40, 215
    frames, cams = synthetic.lambertian_stack(rows, cols, n, seed=17, noise=25.0)
    p = rows * cols
    d_f, d_c = dev(torch, frames), dev(torch, cams)
    coef = torch.empty((6, p), dtype=torch.float64, device="cuda")
    count = torch.empty(p, dtype=torch.int32, device="cuda")
    bad = torch.zeros(1, dtype=torch.int32, device="cuda")
    _capi.call("rti_ptm_fit_frames_masked", d_f.data_ptr(), p, d_c.data_ptr(), n, 0, rows, cols, lo, hi,
               coef.data_ptr(), p, count.data_ptr(), bad.data_ptr(), None)
    torch.cuda.synchronize()
    lx, ly, inten = orc.gather_soa(frames, cams)
    got = coef.cpu().numpy().reshape(6, rows, cols)
    cnt = count.cpu().numpy().reshape(rows, cols)
    n_bad = 0
    checked = 0
    for y in range(rows):
        for x in range(cols):
            m = (inten[y, x] >= lo) & (inten[y, x] <= hi)
            assert cnt[y, x] == m.sum()
            if m.sum() < 6:
                n_bad += 1
                assert np.all(np.isnan(got[:, y, x]))
                continue
            ref = orc.ptm_fit(lx[y, x][m], ly[y, x][m], inten[y, x][m])
            if np.linalg.cond(orc.ptm_design(lx[y, x][m], ly[y, x][m])) < 1e6:
                assert np.abs(got[:, y, x] - ref).max() <= 1e-8 * np.abs(ref).max()
                checked += 1
    assert checked > p // 2 and not (cnt == n).all()          # masks really differ between pixels
    assert bad.item() >= n_bad
    # a full range reproduces the unmasked fit bit for bit
    coef_all, bad_all = gpu_fit_frames(torch, frames, cams)
    _capi.call("rti_ptm_fit_frames_masked", d_f.data_ptr(), p, d_c.data_ptr(), n, 0, rows, cols, 0, 255,
               coef.data_ptr(), p, count.data_ptr(), bad.data_ptr(), None)
    torch.cuda.synchronize()
    assert torch.equal(coef, coef_all) and (count == n).all()


def test_ptm_masked_call_surface(torch_cuda, monkeypatch):
    import rti_b200
    from rti_b200 import analysis, synthetic
    r, n = 8, 40
    frames, cams = synthetic.lambertian_stack(r, r, n, seed=4, noise=20.0)
    monkeypatch.setattr(rti_b200.constants, "ROI_DIAMETER", r)
    res, count = analysis.interpolate_ptm_masked(synthetic.as_records(frames, cams), (10, 245))
    assert res.mode == "ptm" and tuple(count.shape) == (r, r)
    lx, ly, inten = orc.gather_soa(frames, cams)
    m = (inten[3, 5] >= 10) & (inten[3, 5] <= 245)
    ref = orc.ptm_fit(lx[3, 5][m], ly[3, 5][m], inten[3, 5][m])
    got = res.coefficients()[:, 3, 5]
    assert np.abs(got - ref).max() <= 1e-8 * np.abs(ref).max()
    with pytest.raises(rti_b200.RtiSingularError):
        analysis.interpolate_ptm_masked(synthetic.as_records(frames, cams), (254, 255))   # almost nothing valid
