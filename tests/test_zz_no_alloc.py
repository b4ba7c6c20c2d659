"""Allocation-free BN statistics entry point, exercised through the C ABI.

`ldnn_bn_fwd_stats_noalloc` exists for callers that cannot allocate device memory and have no
workspace to hand over — a plug-in or custom op inside a host whose allocator owns the whole GPU, or
a loop that must never reach cudaMalloc/cudaFree because both synchronise the device.  Its partial
sums and ticket live in a fixed 1 MiB array inside the library, so the only thing to verify is that
it computes exactly what `ldnn_bn_fwd_stats` computes with a caller-owned workspace — on two calls
in a row, because the second call finds the first call's leftovers in the shared array.

Collected last on purpose (file name): a `-x` run has already executed every other test.
"""
import ctypes as C

import pytest
import torch

pytestmark = pytest.mark.gpu

EPS, MOM = 1e-5, 0.1


def _vp(t):
    return C.c_void_p(t.data_ptr())


def _outputs(C_):
    dev = "cuda"
    return {"rm": torch.zeros(C_, device=dev), "rv": torch.ones(C_, device=dev),
            "nbt": torch.zeros(1, dtype=torch.int64, device=dev), "sm": torch.zeros(C_, device=dev),
            "si": torch.zeros(C_, device=dev), "sc": torch.zeros(C_, device=dev), "sh": torch.zeros(C_, device=dev)}


@pytest.mark.parametrize("M,C_,chunks", [(64 * 16 * 16, 128, 128), (64 * 2 * 2, 1024, 8), (333, 256, 10)])
def test_noalloc_matches_workspace_entry_point(M, C_, chunks):
    from ldnn_b200.ops import _lib as L
    lib = L.lib()
    g = torch.Generator().manual_seed(M)
    gamma = (torch.rand(C_, generator=g) + 0.5).cuda()
    beta = torch.randn(C_, generator=g).cuda()
    a, b = _outputs(C_), _outputs(C_)
    partial = torch.zeros(chunks * 2 * C_, device="cuda")
    ticket = torch.zeros(1, dtype=torch.int32, device="cuda")
    for call in range(2):                       # two calls in a row, different data each time
        y = (torch.randn(M, C_, generator=g) * (1.5 + call) + 0.25 * call).cuda().bfloat16()
        rc = lib.ldnn_bn_fwd_stats(_vp(y), M, C_, chunks, _vp(partial), _vp(ticket), _vp(gamma), _vp(beta),
                                   _vp(a["rm"]), _vp(a["rv"]), _vp(a["nbt"]), _vp(a["sm"]), _vp(a["si"]),
                                   _vp(a["sc"]), _vp(a["sh"]), C.c_float(EPS), C.c_float(MOM), None)
        assert rc == 0, lib.ldnn_last_error().decode()
        rc = lib.ldnn_bn_fwd_stats_noalloc(_vp(y), M, C_, chunks, _vp(gamma), _vp(beta), _vp(b["rm"]), _vp(b["rv"]),
                                           _vp(b["nbt"]), _vp(b["sm"]), _vp(b["si"]), _vp(b["sc"]), _vp(b["sh"]),
                                           C.c_float(EPS), C.c_float(MOM), None)
        assert rc == 0, lib.ldnn_last_error().decode()
        torch.cuda.synchronize()
        for k in ("sm", "si", "sc", "sh", "rm", "rv"):
            assert torch.allclose(b[k], a[k], atol=1e-4, rtol=1e-3), (call, k)
        assert int(b["nbt"]) == int(a["nbt"]) == call + 1
        # and both agree with plain torch statistics of the same bf16 data
        yf = y.float()
        assert torch.allclose(b["sm"], yf.mean(0), atol=1e-3, rtol=1e-3)
        assert torch.allclose(b["si"], (yf.var(0, unbiased=False) + EPS).rsqrt(), atol=1e-3, rtol=1e-3)


def test_noalloc_refuses_requests_beyond_its_scratch():
    from ldnn_b200.ops import _lib as L
    lib = L.lib()
    C_, chunks, M = 2048, 128, 4096          # 128*2*2048 floats = 2 MiB > the 1 MiB built in
    y = torch.zeros(M, C_, device="cuda", dtype=torch.bfloat16)
    o = _outputs(C_)
    gamma, beta = torch.ones(C_, device="cuda"), torch.zeros(C_, device="cuda")
    rc = lib.ldnn_bn_fwd_stats_noalloc(_vp(y), M, C_, chunks, _vp(gamma), _vp(beta), _vp(o["rm"]), _vp(o["rv"]),
                                       _vp(o["nbt"]), _vp(o["sm"]), _vp(o["si"]), _vp(o["sc"]), _vp(o["sh"]),
                                       C.c_float(EPS), C.c_float(MOM), None)
    assert rc != 0
    assert "exceeds the built-in scratch" in lib.ldnn_last_error().decode()
    torch.cuda.synchronize()
    assert int(o["nbt"]) == 0                # nothing was launched


def test_noalloc_on_a_side_stream():
    """The ticket is cleared and the kernel launched in the order of the stream the caller passes: two calls queued on
    one non-default stream, with the results read only after that stream has drained."""
    from ldnn_b200.ops import _lib as L
    lib = L.lib()
    M, C_, chunks = 4096, 256, 32
    g = torch.Generator().manual_seed(5)
    gamma, beta = (torch.rand(C_, generator=g) + 0.5).cuda(), torch.randn(C_, generator=g).cuda()
    ys = [(torch.randn(M, C_, generator=g) * (1 + i)).cuda().bfloat16() for i in range(2)]
    o = _outputs(C_)
    torch.cuda.synchronize()
    st = torch.cuda.Stream()
    for y in ys:
        rc = lib.ldnn_bn_fwd_stats_noalloc(_vp(y), M, C_, chunks, _vp(gamma), _vp(beta), _vp(o["rm"]), _vp(o["rv"]),
                                           _vp(o["nbt"]), _vp(o["sm"]), _vp(o["si"]), _vp(o["sc"]), _vp(o["sh"]),
                                           C.c_float(EPS), C.c_float(MOM), C.c_void_p(st.cuda_stream))
        assert rc == 0, lib.ldnn_last_error().decode()
    st.synchronize()
    yf = ys[1].float()
    assert int(o["nbt"]) == 2
    assert torch.allclose(o["sm"], yf.mean(0), atol=1e-3, rtol=1e-3)
    assert torch.allclose(o["si"], (yf.var(0, unbiased=False) + EPS).rsqrt(), atol=1e-3, rtol=1e-3)
