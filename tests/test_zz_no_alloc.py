"""chaqsdf_u8_device_static (include/chaqsdf.h): the variant of chaqsdf_u8_device whose workspace and tables live in a
file-scope `static __device__` array inside the library, for callers that cannot call cudaMalloc. No counterpart in the
reference. Called through the C ABI and compared byte for byte (the bar of
every u8 test here: bit-exact) with the existing entry point, on two calls in a row, with the torch caching allocator
watched so that the new entry point is seen to allocate nothing through it. Collected last on purpose (file name)."""
import os
import sys

import numpy as np
import pytest

from conftest import ROOT

sys.path.insert(0, ROOT)
import __graft_entry__ as ge  # noqa: E402

pytestmark = pytest.mark.gpu
torch = pytest.importorskip("torch")


@pytest.fixture(scope="module")
def pkg():
    assert torch.cuda.is_available(), "GPU tests need a CUDA device"
    p = ge.load_package()
    p.lib()
    return p


def shapes(rng, h, w):
    yy, xx = np.mgrid[0:h, 0:w]
    mask = np.zeros((h, w), bool)
    for _ in range(6):
        cy, cx, r = rng.integers(0, h), rng.integers(0, w), rng.integers(2, max(3, min(h, w) // 3))
        m = (yy - cy) ** 2 + (xx - cx) ** 2 <= r * r
        mask = (mask & ~m) if rng.random() < 0.3 else (mask | m)
    return (mask * 255).astype(np.uint8)


def test_static_scratch_matches_workspace_entry_twice(pkg, oracle):
    rng = np.random.default_rng(7)
    # two calls in a row that differ in everything the shared array caches: shape, spread, table kind, symmetric/asymmetric
    cases = [(700, 1000, 40, False), (300, 520, 256, True), (700, 1000, 40, False)]
    for (h, w, spread, asym) in cases:
        img = shapes(rng, h, w)
        d = torch.from_numpy(img).cuda()
        want = pkg.sdf_u8_device(d, w, h, spread=spread, asymmetric=asym)
        out1 = torch.empty((h, w), dtype=torch.uint8, device="cuda")
        out2 = torch.full((h, w), 0x5A, dtype=torch.uint8, device="cuda")
        torch.cuda.synchronize()
        before = torch.cuda.memory_stats()["allocation.all.allocated"]
        pkg.sdf_u8_device_static(d, w, h, out1, spread=spread, asymmetric=asym)
        torch.cuda.synchronize()
        free_before = torch.cuda.mem_get_info()[0]  # cudaMemGetInfo: would drop if the library called cudaMalloc
        pkg.sdf_u8_device_static(d, w, h, out2, spread=spread, asymmetric=asym)  # second call: array holds the first call's contents
        torch.cuda.synchronize()
        assert torch.cuda.mem_get_info()[0] >= free_before
        assert torch.cuda.memory_stats()["allocation.all.allocated"] == before
        assert torch.equal(out1, want), f"first call differs in {(out1 != want).sum().item()} bytes ({h}x{w} s{spread})"
        assert torch.equal(out2, want), f"second call differs in {(out2 != want).sum().item()} bytes ({h}x{w} s{spread})"
        assert np.array_equal(want.cpu().numpy(), oracle.sdf(img, w, h, spread=spread, asymmetric=asym))


def test_static_scratch_on_a_side_stream_and_strided_input(pkg):
    rng = np.random.default_rng(11)
    h, w = 257, 333  # ragged: partial column group, partial block
    ya = np.zeros((h, w, 2), np.uint8)
    ya[..., 1] = shapes(rng, h, w)
    d = torch.from_numpy(ya).cuda()
    want = pkg.sdf_u8_device(d, w, h, stride=2, offset=1, spread=16)
    out = torch.empty((h, w), dtype=torch.uint8, device="cuda")
    torch.cuda.synchronize()
    st = torch.cuda.Stream()
    with torch.cuda.stream(st):
        pkg.sdf_u8_device_static(d, w, h, out, stride=2, offset=1, spread=16, stream=st)
        pkg.sdf_u8_device_static(d, w, h, out, stride=2, offset=1, spread=16, stream=st)
    st.synchronize()
    assert torch.equal(out, want)


def test_static_scratch_rejects_what_does_not_fit(pkg):
    side = 4096
    assert pkg.workspace_bytes(side, side, 64) > pkg.STATIC_SCRATCH_BYTES
    d = torch.zeros((side, side), dtype=torch.uint8, device="cuda")
    out = torch.empty_like(d)
    with pytest.raises(pkg.ChaqSdfError, match="static array holds"):
        pkg.sdf_u8_device_static(d, side, side, out, spread=64)
    small = torch.zeros((64, 64), dtype=torch.uint8, device="cuda")
    with pytest.raises(pkg.ChaqSdfError, match="spread > 256"):
        pkg.sdf_u8_device_static(small, 64, 64, torch.empty_like(small), spread=300)
    # the largest shape the header names fits
    big = torch.zeros((2048, 2048), dtype=torch.uint8, device="cuda")
    big[500:900, 300:1500] = 255
    o = torch.empty_like(big)
    pkg.sdf_u8_device_static(big, 2048, 2048, o, spread=256)
    torch.cuda.synchronize()
    assert torch.equal(o, pkg.sdf_u8_device(big, 2048, 2048, spread=256))
