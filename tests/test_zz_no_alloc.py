"""rti_rbf_fit_soa_static: the allocation-free variant of rti_rbf_fit_soa (workspace in a device array
inside the library; no counterpart in the reference, added for callers that cannot call cudaMalloc).
Checked through the C ABI against rti_rbf_fit_soa on the same inputs, on calls in a row that share
the scratch.  Named test_zz_* so that pytest collects it after every other test file."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

RBF_GRID_RTOL = 1e-8        # the bound test_gpu_parity.py holds the fp64 RBF fit to
STATIC_SCRATCH_BYTES = 32 << 20     # RTI_RBF_STATIC_SCRATCH_BYTES in include/rti_b200.h


@pytest.fixture(scope="module")
def torch_cuda(lib):
    import torch
    assert torch.cuda.is_available()
    torch.cuda.set_device(0)
    return torch


def nodes(p, n, seed):
    rng = np.random.default_rng(seed)
    r = np.sqrt(rng.uniform(0, 0.95, (p, n)))
    th = rng.uniform(0, 2 * np.pi, (p, n))
    return ((r * np.cos(th)).astype(np.float32), (r * np.sin(th)).astype(np.float32),
            rng.integers(0, 256, (p, n)).astype(np.int32))


def upload(torch, arrays):
    return [torch.from_numpy(np.ascontiguousarray(a)).cuda() for a in arrays]


def fit_with_workspace(torch, lib, d_lx, d_ly, d_in):
    from rti_b200 import _capi
    p, n = d_lx.shape
    w = torch.empty((p, n), dtype=torch.float64, device="cuda")
    bad = torch.zeros(1, dtype=torch.int32, device="cuda")
    batch = min(p, 32)
    nbytes = lib.rti_rbf_workspace_bytes(n, batch)
    work = torch.empty(nbytes, dtype=torch.uint8, device="cuda")
    _capi.call("rti_rbf_fit_soa", d_lx.data_ptr(), d_ly.data_ptr(), d_in.data_ptr(), p, n, w.data_ptr(),
               work.data_ptr(), nbytes, batch, bad.data_ptr(), None)
    torch.cuda.synchronize()
    return w.cpu().numpy(), int(bad.item())


def test_static_scratch_fit_equals_workspace_fit_on_calls_in_a_row(torch_cuda, lib):
    torch = torch_cuda
    from rti_b200 import _capi
    # (pixels, nodes): more pixels than fit in the scratch at once (14 in flight at n = 500), a
    # smaller system over what the larger one left behind, padding cases (n = 65: one full tile + 1),
    # and the first shape again
    shapes = [(40, 500), (300, 40), (9, 65), (5, 1000), (40, 500)]
    cases = []
    for i, (p, n) in enumerate(shapes):
        d = upload(torch, nodes(p, n, seed=100 + i))
        want, bad = fit_with_workspace(torch, lib, *d)
        assert bad == 0
        out = torch.full((p, n), float("nan"), dtype=torch.float64, device="cuda")
        cases.append((d, want, out))
    d_bad = torch.zeros(1, dtype=torch.int32, device="cuda")
    torch.cuda.synchronize()

    def run(case):
        (d_lx, d_ly, d_in), _want, out = case
        p, n = d_lx.shape
        _capi.call("rti_rbf_fit_soa_static", d_lx.data_ptr(), d_ly.data_ptr(), d_in.data_ptr(), p, n,
                   out.data_ptr(), d_bad.data_ptr(), None)

    run(cases[0])                       # set-up call: resolves the scratch address on this device
    torch.cuda.synchronize()
    free_before, _ = torch.cuda.mem_get_info()
    for case in cases:                  # back to back on one stream, no synchronisation in between
        run(case)
    torch.cuda.synchronize()
    free_after, _ = torch.cuda.mem_get_info()
    assert free_after == free_before, "the allocation-free entry point changed the device's free memory"
    assert int(d_bad.item()) == 0
    for (_d, want, out) in cases:
        got = out.cpu().numpy()
        assert np.isfinite(got).all()
        scale = np.abs(want).max(axis=1, keepdims=True)
        assert (np.abs(got - want) / scale).max() < RBF_GRID_RTOL


def test_static_scratch_reports_singular_pixels_like_the_workspace_fit(torch_cuda, lib):
    torch = torch_cuda
    from rti_b200 import _capi
    lx, ly, inten = nodes(6, 48, seed=7)
    lx[2, 5], ly[2, 5] = lx[2, 9], ly[2, 9]          # duplicate node: scipy raises LinAlgError
    d = upload(torch, (lx, ly, inten))
    _want, bad_ws = fit_with_workspace(torch, lib, *d)
    out = torch.empty((6, 48), dtype=torch.float64, device="cuda")
    d_bad = torch.zeros(1, dtype=torch.int32, device="cuda")
    _capi.call("rti_rbf_fit_soa_static", d[0].data_ptr(), d[1].data_ptr(), d[2].data_ptr(), 6, 48,
               out.data_ptr(), d_bad.data_ptr(), None)
    torch.cuda.synchronize()
    assert bad_ws >= 1 and int(d_bad.item()) == bad_ws


def test_static_scratch_rejects_sizes_beyond_its_maximum(torch_cuda, lib):
    torch = torch_cuda
    from rti_b200 import _capi
    assert lib.rti_rbf_workspace_bytes(1983, 1) <= STATIC_SCRATCH_BYTES < lib.rti_rbf_workspace_bytes(1984, 1)
    n = 1984
    d = upload(torch, nodes(1, n, seed=3))
    out = torch.zeros((1, n), dtype=torch.float64, device="cuda")
    with pytest.raises(_capi.RtiError) as err:
        _capi.call("rti_rbf_fit_soa_static", d[0].data_ptr(), d[1].data_ptr(), d[2].data_ptr(), 1, n,
                   out.data_ptr(), None, None)
    assert "scratch" in str(err.value) and "1983" in str(err.value)
    torch.cuda.synchronize()
    assert float(out.abs().max().item()) == 0.0         # nothing was launched
