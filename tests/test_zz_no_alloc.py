"""GPU: mcpt_encode_png_static — the output stage without any allocation (static device scratch) — against mcpt_encode_png.

Collected LAST on purpose.  The entry point has no counterpart in the reference; it exists for
callers that cannot call cudaMalloc (CUDA graph capture, serving loops).  Checked through the C ABI: the same bytes as the
existing entry point (tolerance of the PNG tests: none, byte for byte) on two calls in a row — the second on a SMALLER image, so
anything the shared static array kept from the first call would show —, the size limit, and that it needs no mcpt_ctx (it runs
before any context exists when this file is run alone)."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def test_static_scratch_png_equals_the_allocating_entry_point(built_lib):
    import torch
    import montecarlopathtracing_b200 as M
    from montecarlopathtracing_b200 import _capi
    lib = _capi.lib()
    rng = np.random.default_rng(2026)
    cases = []
    for (h, w) in ((1080, 1920), (37, 301)):                       # the largest size the array holds, then a small one
        img = rng.uniform(-0.2, 1.4, size=(h, w, 3)).astype(np.float32)
        img.reshape(-1)[::97] = np.nan
        img.reshape(-1)[5::131] = 50.0
        cases.append((h, w, torch.from_numpy(img).cuda()))
    torch.cuda.synchronize()
    got = []
    for h, w, d in cases:                                          # two calls in a row, no context, no allocation in between
        n = int(lib.mcpt_png_size(w, h))
        assert n + 256 <= _capi.PNG_STATIC_BYTES
        out = np.zeros(n, np.uint8)
        rc = lib.mcpt_encode_png_static(0, C.c_void_p(d.data_ptr()), w, h, out.ctypes.data, n, None)
        assert rc == 0, lib.mcpt_last_error().decode()
        got.append(out)
    ctx = M.Context(0)
    try:
        for (h, w, d), out in zip(cases, got):
            want = ctx.encode_png(d.data_ptr(), w, h)              # the existing entry point (grows a per-context buffer)
            assert len(out) == len(want) and np.array_equal(out, want), "first differing byte at %d" % int(np.nonzero(out != want)[0][0])
        # and once more after other work has used the device: same bytes again (the accumulators are cleared every call)
        h, w, d = cases[0]
        again = M.encode_png_static(d.data_ptr(), w, h)
        assert np.array_equal(again, got[0])
        # device output buffer (what a graph-capturing caller would pass) on a non-default stream
        s = torch.cuda.Stream()
        dev_out = torch.zeros(len(got[1]), dtype=torch.uint8, device="cuda")
        torch.cuda.synchronize()
        h, w, d = cases[1]
        rc = lib.mcpt_encode_png_static(0, C.c_void_p(d.data_ptr()), w, h, C.c_void_p(dev_out.data_ptr()), len(got[1]), C.c_void_p(s.cuda_stream))
        assert rc == 0 and np.array_equal(dev_out.cpu().numpy(), got[1])
    finally:
        ctx.close()


def test_static_scratch_limit_is_enforced(built_lib):
    import torch
    import montecarlopathtracing_b200 as M
    d = torch.zeros(8, device="cuda")
    with pytest.raises(M.McptError, match="more than the static scratch holds"):
        M.encode_png_static(d.data_ptr(), 2048, 2048)              # 12.6 MB > 8 MiB: rejected before anything is launched
    with pytest.raises(M.McptError, match="bad argument"):
        M.encode_png_static(d.data_ptr(), 0, 10)
