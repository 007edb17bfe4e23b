"""ROI extraction: get_ROI(hsv=True) (Utils/utilities.py:302-342) and the
V-channel pick of FeatureMatcher.py:183-184.  CPU tests pin the oracle to the reference's own
outputs (tests/golden/roi.npz, generated with the unmodified ut.get_ROI + cv2); GPU tests hold
rti_roi_extract to the oracle, bit-exact."""
import os

import numpy as np
import pytest

import rti_oracle as orc
from conftest import GOLDEN


@pytest.fixture(scope="module")
def roi_golden():
    return dict(np.load(os.path.join(GOLDEN, "roi.npz")))


def test_oracle_roi_matches_reference_outputs(roi_golden):
    img, diam = roi_golden["img"], int(roi_golden["diameter"])
    for k, corners in enumerate(roi_golden["corners"]):
        want = roi_golden["roi%d" % k]
        got = orc.get_roi_hsv(img, [tuple(c) for c in corners], diam)
        assert got.shape == want.shape
        assert np.array_equal(got, want)
        assert np.array_equal(orc.roi_intensities(img, corners, diam), want[:, :, 2])


def test_roi_window_quirks():
    # Utils/utilities.py:320-327: negative lower bound -> 1, upper bound past the edge -> size - 1
    assert orc.roi_bounds((90, 120), [(2, 3), (20, 25)], 40) == (1, 31, 1, 34)
    assert orc.roi_bounds((90, 120), [(100, 70), (119, 89)], 40) == (90, 119, 60, 89)
    # centre uses Python round (half to even): (41+80)/2 = 60.5 -> 80 - 19.5 = 60.5 -> 60
    assert orc.corners_center([(41, 30), (80, 61)], 90, 120) == (60, 46)


def test_host_window_equals_oracle(roi_golden):
    from rti_b200 import roi
    for corners in roi_golden["corners"]:
        for diam in (40, 41, 7, 200):
            assert roi.roi_window((90, 120), corners, diam) == orc.roi_bounds((90, 120), corners, diam)


def test_bgr2hsv_known_values():
    px = np.array([[[0, 0, 0], [255, 255, 255], [255, 0, 0], [0, 255, 0], [0, 0, 255], [10, 200, 90]]], np.uint8)
    hsv = orc.bgr2hsv_u8(px)[0]
    assert hsv[0].tolist() == [0, 0, 0] and hsv[1].tolist() == [0, 0, 255]
    assert hsv[2].tolist() == [120, 255, 255] and hsv[3].tolist() == [60, 255, 255] and hsv[4].tolist() == [0, 255, 255]
    assert hsv[5, 2] == 200


# ------------------------------------------------------------------------------- GPU
@pytest.fixture(scope="module")
def torch_cuda(lib):
    import torch
    assert torch.cuda.is_available()
    torch.cuda.set_device(0)
    return torch


@pytest.mark.gpu
def test_gpu_get_roi_matches_reference_golden(torch_cuda, roi_golden):
    import rti_b200
    from rti_b200 import roi
    img, diam = roi_golden["img"], int(roi_golden["diameter"])
    old = rti_b200.constants.ROI_DIAMETER
    rti_b200.constants.ROI_DIAMETER = diam
    try:
        for k, corners in enumerate(roi_golden["corners"]):
            got = roi.get_ROI(img, [tuple(c) for c in corners], hsv=True)
            assert np.array_equal(got, roi_golden["roi%d" % k]), "case %d" % k
    finally:
        rti_b200.constants.ROI_DIAMETER = old


@pytest.mark.gpu
def test_gpu_bgr2hsv_exhaustive(torch_cuda):
    """All 2^24 colours, as a 4096 x 4096 frame cropped at an odd offset."""
    torch = torch_cuda
    from rti_b200 import roi
    a = np.arange(1 << 24, dtype=np.uint32)
    cube = np.stack([a & 255, (a >> 8) & 255, (a >> 16) & 255], -1).astype(np.uint8).reshape(4096, 4096, 3)
    d = torch.from_numpy(cube).cuda()
    got = roi.extract_device(d, (0, 4096, 0, 4096), hsv=True).cpu().numpy()
    assert np.array_equal(got, orc.bgr2hsv_u8(cube))
    win = (3, 4094, 5, 1001)
    got = roi.extract_device(d, win, hsv=True).cpu().numpy()
    assert np.array_equal(got, orc.bgr2hsv_u8(cube[5:1001, 3:4094]))


@pytest.mark.gpu
@pytest.mark.parametrize("w,x_min,cols", [(120, 0, 120), (121, 1, 119), (1919, 759, 400), (1919, 760, 401),
                                          (1300, 5, 1027), (64, 63, 1), (2300, 11, 2051)])
def test_gpu_frames_to_stack_every_alignment(torch_cuda, w, x_min, cols):
    """Batched V-channel stack (FeatureMatcher.py:183-184) for row starts / widths of every
    alignment, including runs longer than one CTA tile and 1-pixel ROIs."""
    torch = torch_cuda
    from rti_b200 import roi
    rng = np.random.default_rng(w * 7 + cols)
    h, f = 37, 3
    frames = rng.integers(0, 256, (f, h, w, 3), dtype=np.uint8)
    win = (x_min, x_min + cols, 2, 33)
    d = torch.from_numpy(frames).cuda()
    v = roi.extract_device(d, win, hsv=False).cpu().numpy()
    want = frames[:, 2:33, x_min:x_min + cols].max(axis=-1)
    assert v.shape == want.shape and np.array_equal(v, want)
    hsv = roi.extract_device(d, win, hsv=True).cpu().numpy()
    assert np.array_equal(hsv, orc.bgr2hsv_u8(frames[:, 2:33, x_min:x_min + cols]))


@pytest.mark.gpu
def test_gpu_roi_feeds_the_fit(torch_cuda):
    """BGR frames -> rti_roi_extract -> FrameStack -> PTM fit equals the fit on the oracle's
    ROI intensities: the f4 stage plugs into the path without a host round trip."""
    torch = torch_cuda
    from rti_b200 import roi, analysis
    rng = np.random.default_rng(5)
    n, h, w, diam = 24, 40, 52, 12
    frames = rng.integers(0, 256, (n, h, w, 3), dtype=np.uint8)
    corners = [(20, 14), (33, 14), (33, 27), (20, 27)]
    cams = np.column_stack([rng.uniform(-10, 22, n), rng.uniform(-10, 22, n), rng.uniform(9, 27, n)])
    win = orc.roi_bounds((h, w), corners, diam)
    stack = roi.extract_device(torch.from_numpy(frames).cuda(), roi.roi_window((h, w), corners, diam), hsv=False)
    want = np.stack([orc.roi_intensities(fr, corners, diam) for fr in frames])
    assert np.array_equal(stack.cpu().numpy(), want)
    fs = analysis.FrameStack.from_device(stack, torch.from_numpy(cams).cuda())
    coef_dev, bad = analysis._fit_ptm_from_stack(fs)
    assert int(bad.item()) == 0
    lx, ly, inten = orc.gather_soa(want, cams)
    coef_ref = orc.ptm_fit_batch(lx, ly, inten)
    coef = coef_dev.cpu().numpy().reshape(6, want.shape[1], want.shape[2]).transpose(1, 2, 0)
    assert np.abs(coef - coef_ref).max() <= 1e-9 * np.abs(coef_ref).max()
    assert win == roi.roi_window((h, w), corners, diam)


@pytest.mark.gpu
def test_gpu_roi_argument_errors(torch_cuda):
    torch = torch_cuda
    from rti_b200 import _capi
    d = torch.zeros((10, 10, 3), dtype=torch.uint8, device="cuda")
    o = torch.zeros((10, 10, 3), dtype=torch.uint8, device="cuda")
    with pytest.raises(_capi.RtiError):
        _capi.call("rti_roi_extract", d.data_ptr(), 300, 1, 10, 10, 5, 0, 10, 10, 1, o.data_ptr(), 300, None)
    with pytest.raises(_capi.RtiError):
        _capi.call("rti_roi_extract", None, 300, 1, 10, 10, 0, 0, 10, 10, 1, o.data_ptr(), 300, None)


@pytest.mark.gpu
def test_gpu_roi_stays_inside_its_buffers(torch_cuda):
    """Guard bands (no compute-sanitizer on the pool): the frame is allocated exactly, the ROI touches its
    last byte, and nothing outside the output is written."""
    torch = torch_cuda
    from rti_b200 import _capi
    rng = np.random.default_rng(3)
    h, w = 21, 37
    frame = rng.integers(0, 256, (h, w, 3), dtype=np.uint8)
    d = torch.from_numpy(frame).cuda()
    held = torch.zeros(h * w * 3 + 64, dtype=torch.uint8, device="cuda")     # the same frame at an odd address
    held[13:13 + h * w * 3] = d.reshape(-1)
    for hsv, bpp, src in ((0, 1, d.data_ptr()), (1, 3, d.data_ptr()), (0, 1, held.data_ptr() + 13),
                          (1, 3, held.data_ptr() + 13)):
        rows, cols, x0, y0 = 9, 30, 7, 12                     # ends at the frame's last pixel
        guard = 256
        out = torch.full((guard + rows * cols * bpp + guard,), 0xAB, dtype=torch.uint8, device="cuda")
        _capi.call("rti_roi_extract", src, h * w * 3, 1, h, w, x0, y0, rows, cols, hsv,
                   out.data_ptr() + guard, rows * cols * bpp, None)
        torch.cuda.synchronize()
        o = out.cpu().numpy()
        assert np.all(o[:guard] == 0xAB) and np.all(o[-guard:] == 0xAB)
        want = orc.bgr2hsv_u8(frame[y0:y0 + rows, x0:x0 + cols])
        want = want if hsv else want[:, :, 2]
        assert np.array_equal(o[guard:-guard].reshape(want.shape), want)
