#!/usr/bin/env python
"""ROI extraction alone (development aid): V and HSV at 40 frames of 3840x2160, ROI 2048^2, and at
the reference's default ROI 400^2 x 300 frames of 1920x1080; checks one frame against torch."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

from rti_b200 import _capi  # noqa: E402

torch.cuda.set_device(0)
S = torch.cuda.current_stream().cuda_stream
for fh, fw, nf, roi in ((2160, 3840, 40, 2048), (1080, 1920, 300, 400)):
    bgr = torch.randint(0, 256, (nf, fh, fw, 3), dtype=torch.uint8, device="cuda")
    x0, y0 = (fw - roi) // 2 + 1, (fh - roi) // 2 + 1
    for hsv in (0, 1):
        bpp = 3 if hsv else 1
        o = torch.empty((nf, roi, roi, bpp), dtype=torch.uint8, device="cuda")
        run = lambda: _capi.call("rti_roi_extract", bgr.data_ptr(), fh * fw * 3, nf, fh, fw, x0, y0, roi, roi, hsv,
                                 o.data_ptr(), roi * roi * bpp, S)
        for _ in range(5):
            run()
        ev = [torch.cuda.Event(enable_timing=True) for _ in range(2)]
        ev[0].record()
        for _ in range(20):
            run()
        ev[1].record()
        torch.cuda.synchronize()
        ms = ev[0].elapsed_time(ev[1]) / 20
        ok = True
        if not hsv:
            ok = bool(torch.equal(o[3, :, :, 0], bgr[3, y0:y0 + roi, x0:x0 + roi].amax(dim=2)))
        print("roi %4d x %3d frames hsv=%d  %.4f ms  %.0f GB/s  ok=%s" % (roi, nf, hsv, ms, (3 + bpp) * nf * roi * roi / ms / 1e6, ok))
