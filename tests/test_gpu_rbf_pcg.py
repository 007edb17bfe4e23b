"""Tile-preconditioned linear-RBF fit (rti_rbf_fit_pcg, csrc/rbf_pcg.cu) against the oracle's
scipy restatement (analysis.py:249-260): node residuals, weights' grid values within 5e-4 of the
grid maximum (north_star: 1e-3), partial tiles, iteration counts, the unconverged-pixel report."""
import numpy as np
import pytest

import rti_oracle as orc

pytestmark = pytest.mark.gpu

PCG_GRID_RTOL = 5e-4


@pytest.fixture(scope="module")
def torch_cuda(lib):
    import torch
    assert torch.cuda.is_available()
    torch.cuda.set_device(0)
    return torch


def slab(rows, cols, n, seed, roi=1024, row0=500, col0=480, lambert=False):
    """A rows x cols window of a roi x roi ROI lit from n camera positions (per-pixel light vectors
    of the reference's geometry: ~1e-3 change per pixel step at roi = 1024)."""
    from rti_b200 import synthetic
    rng = np.random.default_rng(seed)
    cams = synthetic.camera_positions(roi, roi, n, rng)
    if lambert:
        frames, _ = synthetic.lambertian_stack(rows, cols, n, seed=seed, row0=row0, rows_total=roi)
    else:
        frames = rng.integers(0, 256, (n, rows, cols), dtype=np.uint8)          # white noise: the worst case
    return orc.gather_soa(frames, cams, row0=row0, col0=col0)


def fit_pcg(torch, lx, ly, inten, max_iters=80, tol=0.01, tiles_in_flight=None):
    from rti_b200 import _capi
    lib = _capi.load()
    rows, cols, n = lx.shape
    d = [torch.from_numpy(np.ascontiguousarray(a.reshape(rows * cols, n))).cuda() for a in (lx, ly, inten)]
    tiles = ((rows + 15) // 16) * ((cols + 15) // 16)
    nbytes = lib.rti_rbf_pcg_workspace_bytes(n, tiles if tiles_in_flight is None else tiles_in_flight)
    assert nbytes > 0
    work = torch.empty(nbytes, dtype=torch.uint8, device="cuda")
    w = torch.full((rows * cols, n), float("nan"), dtype=torch.float64, device="cuda")
    iters = torch.full((rows * cols,), -1, dtype=torch.int32, device="cuda")
    bad = torch.zeros(2, dtype=torch.int32, device="cuda")
    _capi.call("rti_rbf_fit_pcg", d[0].data_ptr(), d[1].data_ptr(), d[2].data_ptr(), rows, cols, n, w.data_ptr(),
               work.data_ptr(), nbytes, max_iters, tol, iters.data_ptr(), bad.data_ptr(), None)
    torch.cuda.synchronize()
    return w.cpu().numpy().reshape(rows, cols, n), iters.cpu().numpy().reshape(rows, cols), bad.cpu().numpy()


def check_pixels(lx, ly, inten, w, pixels, node_atol):
    ax = orc.light_axis()
    worst = 0.0
    for (y, x) in pixels:
        a = orc.rbf_matrix(lx[y, x], ly[y, x])
        assert np.max(np.abs(a @ w[y, x] - inten[y, x])) <= node_atol, (y, x)
        ref = orc.rbf_grid(lx[y, x], ly[y, x], orc.rbf_fit(lx[y, x], ly[y, x], inten[y, x]), ax)
        got = orc.rbf_grid(lx[y, x], ly[y, x], w[y, x], ax)
        worst = max(worst, np.max(np.abs(got - ref)) / np.max(np.abs(ref)))
    assert worst <= PCG_GRID_RTOL, worst
    return worst


@pytest.mark.parametrize("rows,cols,n,batch", [(20, 37, 40, None), (16, 16, 130, None), (33, 16, 300, 2),
                                               (17, 20, 500, None), (16, 16, 3, None),
                                               (5, 18, 65, None), (18, 5, 129, 1), (3, 3, 257, None)])
def test_pcg_fit_vs_oracle(torch_cuda, rows, cols, n, batch):
    lx, ly, inten = slab(rows, cols, n, seed=rows + n)
    w, iters, bad = fit_pcg(torch_cuda, lx, ly, inten, tiles_in_flight=batch)
    assert bad[0] == 0 and bad[1] == np.count_nonzero(iters < 0) and bad[1] <= 0.03 * iters.size
    assert np.all(np.isfinite(w))
    assert np.abs(iters).min() >= 1 and np.abs(iters).max() < 80
    rng = np.random.default_rng(1)
    pixels = {(0, 0), (rows - 1, cols - 1), (min(8, rows - 1), min(8, cols - 1))}
    while len(pixels) < min(14, rows * cols):
        pixels.add((int(rng.integers(rows)), int(rng.integers(cols))))
    pixels = [yx for yx in sorted(pixels) if iters[yx] > 0]      # (flagged pixels belong to the fp64 kernels)
    check_pixels(lx, ly, inten, w, pixels, node_atol=0.12)
    # the tile-centre pixel is solved by its own inverse: a couple of steps
    cy, cx = min(8, rows - 1), min(8, cols - 1)
    assert iters[cy, cx] <= 3


def test_pcg_n1000_lambertian(torch_cuda):
    """Config 4's light count on smooth (Lambertian) intensities."""
    lx, ly, inten = slab(16, 16, 1000, seed=5, lambert=True)
    w, iters, bad = fit_pcg(torch_cuda, lx, ly, inten, max_iters=120)
    assert bad.tolist() == [0, 0]
    check_pixels(lx, ly, inten, w, [(0, 0), (8, 8), (15, 15), (3, 12)], node_atol=0.06)


def test_pcg_reports_unconverged_pixels(torch_cuda):
    lx, ly, inten = slab(16, 16, 130, seed=9)
    _, iters, bad = fit_pcg(torch_cuda, lx, ly, inten, max_iters=2)
    assert bad[0] == 0 and bad[1] > 0
    assert bad[1] == np.count_nonzero(iters < 0) and np.all(np.abs(iters) <= 2)


def _py_fit(torch, lx, ly, inten, **kw):
    from rti_b200 import analysis
    rows, cols, n = lx.shape
    d = [torch.from_numpy(np.ascontiguousarray(a.reshape(rows * cols, n))).cuda() for a in (lx, ly, inten)]
    bad = torch.zeros(1, dtype=torch.int32, device="cuda")
    w, info = analysis.rbf_fit_device_pcg(d[0], d[1], d[2], rows, cols, bad, **kw)
    torch.cuda.synchronize()
    return w.cpu().numpy().reshape(rows, cols, n), info, int(bad.item())


def test_pcg_unrelated_neighbours_are_refitted_directly(torch_cuda):
    """Data whose neighbouring pixels do NOT have similar nodes (arbitrary arrays handed to
    interpolate_intensities): the tile-centre preconditioner is useless, the host layer refits what did
    not converge with the direct fp64 kernel, and every pixel still matches the oracle."""
    rng = np.random.default_rng(3)
    rows, cols, n = 16, 20, 60
    r = np.sqrt(rng.uniform(0, 0.95, (rows, cols, n)))
    th = rng.uniform(0, 2 * np.pi, (rows, cols, n))
    lx, ly = (r * np.cos(th)).astype(np.float32), (r * np.sin(th)).astype(np.float32)
    inten = rng.integers(0, 256, (rows, cols, n)).astype(np.int32)
    w, info, bad = _py_fit(torch_cuda, lx, ly, inten, max_iters=12)
    assert bad == 0 and info["refit"] > rows * cols // 2
    check_pixels(lx, ly, inten, w, [(0, 0), (8, 8), (15, 19), (3, 17), (12, 1)], node_atol=0.06)


def test_pcg_singular_tile_centre_and_singular_pixel(torch_cuda):
    """Duplicate nodes (scipy: LinAlgError) at a tile CENTRE leave that tile without a preconditioner: its
    pixels are refitted directly and only the genuinely singular pixels are counted."""
    lx, ly, inten = slab(16, 32, 50, seed=4)
    lx, ly = lx.copy(), ly.copy()
    lx[8, 8, 7], ly[8, 8, 7] = lx[8, 8, 3], ly[8, 8, 3]              # centre of tile 0
    lx[2, 20, 9], ly[2, 20, 9] = lx[2, 20, 0], ly[2, 20, 0]          # an ordinary pixel of tile 1
    w, info, bad = _py_fit(torch_cuda, lx, ly, inten)
    assert info["singular_tiles"] == 1 and bad == 2
    assert info["refit"] >= 256
    ok = [(0, 0), (8, 9), (15, 15), (0, 31), (9, 24)]
    check_pixels(lx, ly, inten, w, ok, node_atol=0.06)


def test_pcg_largest_supported_n(torch_cuda):
    lx, ly, inten = slab(16, 16, 1025, seed=6, lambert=True)
    w, iters, bad = fit_pcg(torch_cuda, lx, ly, inten, max_iters=120)
    assert bad.tolist() == [0, 0]
    check_pixels(lx, ly, inten, w, [(0, 0), (9, 7)], node_atol=0.06)


def test_pcg_full_width_slab_interpolates_every_pixel(torch_cuda):
    """Size-independent property at the bench geometry (1024-wide ROI rows, N = 500, white-noise intensities):
    the linear-RBF interpolant reproduces the node values (scipy's defining property, smooth = 0) for EVERY
    pixel of a 16 x 1024 slab: max |A w - d| in fp64 (torch) from the pixel's own nodes.  The iteration solves
    the float32-rounded matrix, so the C entry reports the few pixels whose weights are too large for that
    (node pairs ~1e-5 apart: |w|_2 > 8e5) like unconverged ones; the host layer refits those in fp64.  Also:
    the worst pixel's table against the oracle, and a different tiling of the same pixels."""
    from rti_b200 import _capi, analysis, synthetic
    torch = torch_cuda
    rows, cols, n = 16, 1024, 500
    rng = np.random.default_rng(8)
    cams = synthetic.camera_positions(1024, 1024, n, rng)
    frames = rng.integers(0, 256, (n, rows, cols), dtype=np.uint8)
    d_f, d_c = torch.from_numpy(frames).cuda(), torch.from_numpy(cams).cuda()
    px = rows * cols
    lx = torch.empty((px, n), dtype=torch.float32, device="cuda")
    ly = torch.empty_like(lx)
    inten = torch.empty((px, n), dtype=torch.int32, device="cuda")
    _capi.call("rti_gather_soa", d_f.data_ptr(), px, d_c.data_ptr(), n, 512, rows, cols,
               lx.data_ptr(), ly.data_ptr(), inten.data_ptr(), None)
    lib = _capi.load()

    def residuals(w):
        out = torch.empty(px, dtype=torch.float64, device="cuda")
        for p0 in range(0, px, 512):
            x = lx[p0:p0 + 512].double()
            y = ly[p0:p0 + 512].double()
            a = torch.sqrt((x[:, :, None] - x[:, None, :]) ** 2 + (y[:, :, None] - y[:, None, :]) ** 2)
            out[p0:p0 + 512] = (torch.bmm(a, w[p0:p0 + 512, :, None])[:, :, 0] - inten[p0:p0 + 512].double()).abs().amax(dim=1)
        return out

    # (1) the C entry: flagged pixels are few, every unflagged pixel interpolates
    tiles = (rows // 16) * (cols // 16)
    nbytes = lib.rti_rbf_pcg_workspace_bytes(n, tiles)
    work = torch.empty(nbytes, dtype=torch.uint8, device="cuda")
    w = torch.empty((px, n), dtype=torch.float64, device="cuda")
    iters = torch.empty(px, dtype=torch.int32, device="cuda")
    bad = torch.zeros(2, dtype=torch.int32, device="cuda")
    _capi.call("rti_rbf_fit_pcg", lx.data_ptr(), ly.data_ptr(), inten.data_ptr(), rows, cols, n, w.data_ptr(),
               work.data_ptr(), nbytes, 48, 0.01, iters.data_ptr(), bad.data_ptr(), None)
    torch.cuda.synchronize()
    flagged = iters < 0
    assert bad[0].item() == 0 and bad[1].item() == int(flagged.sum().item())
    assert 0 < int(flagged.sum().item()) < 0.03 * px
    res = residuals(w)
    assert res[~flagged].max().item() <= 0.2
    assert res.max().item() > 0.2                       # the flag is not vacuous on this data
    # (2) the host layer: flagged pixels refitted in fp64, every pixel interpolates
    hbad = torch.zeros(1, dtype=torch.int32, device="cuda")
    w2, info = analysis.rbf_fit_device_pcg(lx, ly, inten, rows, cols, hbad, max_iters=48)
    assert hbad.item() == 0 and info["refit"] == int(flagged.sum().item())
    res2 = residuals(w2)
    assert res2.max().item() <= 0.2
    assert res2[info["fp64_idx"]].max().item() <= 1e-6
    # (3) the pixel that was worst before the refit: its table (fp64 evaluation of flagged pixels) vs the oracle
    q = int(res.argmax().item())
    ax = orc.light_axis()
    tab = torch.empty((100 * 100, px), dtype=torch.int32, device="cuda")
    tb = torch.zeros(1, dtype=torch.int32, device="cuda")
    analysis.rbf_table_device(lx, ly, w2, torch.from_numpy(ax).cuda(), 100, tab, tb, fp64_idx=info["fp64_idx"])
    l1, l2, dd = lx[q].cpu().numpy(), ly[q].cpu().numpy(), inten[q].cpu().numpy()
    ref = orc.rbf_grid(l1, l2, orc.rbf_fit(l1, l2, dd), ax)
    t = tab[:, q].cpu().numpy().reshape(100, 100)
    diff = t != np.trunc(ref).astype(np.int32)
    assert np.all(np.abs(ref[diff] - np.rint(ref[diff])) < 1e-5 * np.max(np.abs(ref)))
    # (4) a one-row slab is a different tiling of the same pixels (16 x 1 tiles): same interpolant
    w1, _ = analysis.rbf_fit_device_pcg(lx[:cols], ly[:cols], inten[:cols], 1, cols, hbad, max_iters=48)
    scale = w2[:cols].abs().amax(dim=1)
    assert ((w1 - w2[:cols]).abs().amax(dim=1) / scale).max().item() < 5e-3
