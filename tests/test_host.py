"""CPU-only checks of the host side: the C-ABI library loads and exports every symbol include/chaqsdf.h
declares; the pass-2 algorithm (the same envelope_core.h / tables.h the kernels compile) emulated on the host
matches the oracle; the row-band halo exchange works with world_size 2 and 3 over gloo."""
import ctypes as C
import os
import re
import subprocess
import sys

import numpy as np
import pytest

from conftest import ROOT

sys.path.insert(0, ROOT)
import __graft_entry__ as ge  # noqa: E402


@pytest.fixture(scope="session")
def pkg():
    ge.build()
    return ge.load_package()


def test_library_exports_every_declared_symbol(pkg):
    hdr = open(os.path.join(ROOT, "include", "chaqsdf.h")).read()
    declared = set(re.findall(r"\b(chaqsdf_\w+|dist_transform_2d)\s*\(", hdr))
    assert declared == set(pkg.SYMBOLS), declared ^ set(pkg.SYMBOLS)
    L = pkg.lib()
    for s in declared:
        assert hasattr(L, s), s
    assert L.chaqsdf_version() == 110


def test_no_gpu_means_error_not_fallback(pkg):
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    img = np.zeros((8, 8), np.uint8)
    with pytest.raises(pkg.ChaqSdfError):
        pkg.sdf_u8(img, 8, 8, spread=4)
    f = np.zeros((8, 8), np.float32)
    with pytest.raises(pkg.ChaqSdfError):
        pkg.dist_transform_2d(f)


def test_product_does_not_touch_the_oracle():
    for dp, _, fs in os.walk(os.path.join(ROOT, "chaq-sdfgen_b200")):
        for f in fs:
            if f.endswith((".py", ".cu", ".cuh", ".h", ".cpp", ".c")) or f == "Makefile":
                txt = open(os.path.join(dp, f), errors="ignore").read()
                # mentions in comments are fine; loading, importing, linking or including is not
                assert not re.search(r"oracle/|liboracle|libchaqref|import\s+oracle|from\s+oracle|oracle\.py|-loracle", txt), \
                    os.path.join(dp, f)


@pytest.fixture(scope="session")
def emul(tmp_path_factory):
    so = str(tmp_path_factory.mktemp("emul") / "libemul.so")
    subprocess.run(["/usr/bin/g++", "-O2", "-std=c++17", "-ffp-contract=off", "-fPIC", "-shared", "-o", so,
                    os.path.join(ROOT, "tests", "emul", "emul.cpp")], check=True)
    L = C.CDLL(so)
    u8p = np.ctypeslib.ndpointer(np.uint8, flags="C")
    i32p = np.ctypeslib.ndpointer(np.int32, flags="C")
    L.emul_column_pass.argtypes = [u8p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, i32p, u8p, C.c_void_p]
    L.emul_quantise.argtypes = [i32p, u8p, C.c_size_t, C.c_size_t, C.c_int, u8p]
    L.emul_ref_byte.argtypes = [C.c_int, C.c_longlong, C.c_size_t, C.c_int]
    L.emul_ref_byte.restype = C.c_uint8
    return L


def _emul_u8(L, mask, spread, asym, band, wide=0, cap=None):
    h, w = mask.shape
    d2 = np.empty((h, w), np.int32)
    cls = np.empty((h, w), np.uint8)
    rc = L.emul_column_pass(np.ascontiguousarray(mask), w, h, spread + 1 if cap is None else cap, band, wide, d2, cls, None)
    assert rc == 0, f"out-of-bounds ring/row access in the emulated column pass (rc={rc})"
    out = np.empty((h, w), np.uint8)
    L.emul_quantise(d2, cls, h * w, spread, int(asym), out)
    return out, d2, cls


def test_emulated_column_pass_matches_oracle(emul, oracle):
    rng = np.random.default_rng(5)
    for it in range(120):
        h, w = int(rng.integers(1, 70)), int(rng.integers(1, 40))
        dens = rng.choice([0.0, 0.01, 0.05, 0.3, 0.5, 0.9, 1.0])
        mask = (rng.random((h, w)) < dens).astype(np.uint8)
        if it % 3 == 0:
            yy, xx = np.mgrid[0:h, 0:w]
            mask = np.zeros((h, w), np.uint8)
            for _ in range(3):
                cy, cx, r = rng.integers(0, h), rng.integers(0, w), rng.integers(1, 12)
                mask |= ((yy - cy) ** 2 + (xx - cx) ** 2 <= r * r).astype(np.uint8)
        img = (mask * 255).astype(np.uint8)
        for spread in (1, 2, 5, 16, 100):
            for band in (0, 7, 33):
                want = oracle.sdf(img, w, h, spread=spread, asymmetric=bool(it & 1))
                got, _, cls = _emul_u8(emul, mask, spread, bool(it & 1), band)
                assert np.array_equal(got, want) and np.array_equal(cls, mask), (it, h, w, spread, band)
        # unwindowed: exact squared distance to the other class
        _, d2, _ = _emul_u8(emul, mask, 4, False, 13, wide=1, cap=1 << 20)
        opp = np.where(mask == 1, oracle.edt_d2(mask, 0), oracle.edt_d2(mask, 1))
        assert np.array_equal(d2, opp)


def test_emulated_discs_and_noise(emul, oracle):
    img = oracle.syn_discs(384)
    mask = (img > 127).astype(np.uint8)
    for spread, band in ((16, 0), (64, 100), (300, 0)):
        got, _, _ = _emul_u8(emul, mask, spread, False, band)
        assert np.array_equal(got, oracle.sdf(img, 384, 384, spread=spread))
    big = oracle.syn_discs(1024)  # long stretches without sites: stale stack entries must be forgotten
    bmask = (big > 127).astype(np.uint8)
    for spread in (2, 33, 100):
        got, _, _ = _emul_u8(emul, bmask, spread, False, 0)
        assert np.array_equal(got, oracle.sdf(big, 1024, 1024, spread=spread))
    noise = oracle.syn_noise(200)
    got, _, _ = _emul_u8(emul, (noise > 127).astype(np.uint8), 8, True, 50)
    assert np.array_equal(got, oracle.sdf(noise, 200, 200, spread=8, asymmetric=True))


def test_emulated_structured_shapes(emul, oracle):
    """Shapes much larger than the spread: exercises the window ("forgetting") logic of envelope_core.h."""
    rng = np.random.default_rng(2)
    for it in range(80):
        h, w = int(rng.integers(20, 260)), int(rng.integers(8, 120))
        yy, xx = np.mgrid[0:h, 0:w]
        mask = np.zeros((h, w), np.uint8)
        for _ in range(int(rng.integers(1, 8))):
            cy, cx, r = rng.integers(-20, h + 20), rng.integers(-20, w + 20), rng.integers(1, 90)
            kind = it % 4
            if kind == 0:
                m = (yy - cy) ** 2 + (xx - cx) ** 2 <= r * r
            elif kind == 1:
                m = abs(yy - cy) + abs(xx - cx) <= r
            elif kind == 2:
                m = abs((yy - cy) * int(rng.integers(1, 4)) - (xx - cx) * int(rng.integers(1, 4))) <= r // 3
            else:
                m = (yy - cy) ** 2 * int(rng.integers(1, 5)) + (xx - cx) ** 2 <= r * r
            mask = (mask & ~m.astype(np.uint8)) if rng.random() < 0.3 else (mask | m.astype(np.uint8))
        img = (mask * 255).astype(np.uint8)
        for spread in (int(rng.integers(1, 6)), int(rng.integers(6, 20)), int(rng.integers(20, 70))):
            for band in (0, int(rng.integers(1, 80))):
                got, _, _ = _emul_u8(emul, mask, spread, False, band)
                assert np.array_equal(got, oracle.sdf(img, w, h, spread=spread)), (it, h, w, spread, band)


def test_quantiser_thresholds_equal_float_formula(emul, oracle):
    # every integer D2 up to beyond the cap, both classes, several spreads: table look-up == float pipeline
    for spread, asym in ((1, 0), (7, 0), (64, 0), (64, 1), (100, 1), (256, 0), (1024, 0)):
        top = (spread + 3) ** 2
        step = 1 if top < 200000 else 7
        d2 = np.arange(0, top, step, dtype=np.int32)
        d2 = np.concatenate([d2, np.array([2**31 - 1], np.int32)])
        for cls in (0, 1):
            c = np.full(d2.shape, cls, np.uint8)
            out = np.empty(d2.shape, np.uint8)
            emul.emul_quantise(d2, c, d2.size, spread, asym, out)
            want = np.array([emul.emul_ref_byte(cls, int(v), spread, asym) for v in d2[:-1]], np.uint8)
            assert np.array_equal(out[:-1], want), (spread, asym, cls)
            assert out[-1] == want[-1]  # "nothing in reach" saturates like the cap
            # and the formula itself is the oracle's (sdfgen.c:98-106, :75-96)
            for v in (1, 2, spread * spread, spread * spread + 1):
                sd = oracle.L.oracle_signed_from_d2(v, 0) if cls else oracle.L.oracle_signed_from_d2(0, v)
                assert oracle.L.oracle_byte_from_signed(sd, spread, asym) == emul.emul_ref_byte(cls, v, spread, asym)


GLOO_WORKER = r'''
import os, sys, torch, torch.distributed as dist
sys.path.insert(0, sys.argv[1])
import __graft_entry__ as ge
pkg = ge.load_package()
rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
dist.init_process_group("gloo", rank=rank, world_size=world)
for (h, w, spread) in ((64, 5, 7), (10, 3, 4), (37, 8, 100)):
    r0, r1, ht, hb, recvs, sends = pkg.halo_plan(h, world, rank, spread)
    buf = torch.full((ht + (r1 - r0) + hb, w), -1, dtype=torch.int16)
    buf[ht:ht + r1 - r0] = torch.arange(r0, r1, dtype=torch.int16)[:, None]
    pkg.exchange_halos(buf, h, world, rank, spread)
    want = torch.arange(r0 - ht, r1 + hb, dtype=torch.int16)[:, None].expand(-1, w)
    assert torch.equal(buf, want), (rank, h, spread, buf[:, 0].tolist())
    assert ht == min(spread, r0) and hb == min(spread, h - r1)
# the transposed-band variant's host side (sdf_u8_sharded_t): the INPUT bands live in a two-slot buffer laid out like the
# halo path's intermediates; the band goes into tband_input_view, the context rows arrive through the same exchange
img = (torch.arange(53 * 6, dtype=torch.int32) % 251).to(torch.uint8).reshape(53, 6)
for spread in (5, 9, 40):
    state = {}
    pkg.tband_setup(state, 6, 53, spread, rank, world, None, torch.device("cpu"))
    assert state["halo"] == "p2p"                       # no peer memory on the CPU: gloo point to point
    r0, r1, ht, hb, _, _ = state["plans"][rank]
    for call in range(3):                               # slots alternate by call
        slot = state["step"] & 1
        view = pkg.tband_input_view(state, rank)
        assert view.shape == (r1 - r0, 6) and view.data_ptr() == state["sym"][slot][ht:ht + r1 - r0].data_ptr()
        view.copy_(img[r0:r1])
        ext = state["sym"][slot][:ht + (r1 - r0) + hb]
        pkg.exchange_halos(ext, 53, world, rank, spread)
        assert torch.equal(ext, img[r0 - ht:r1 + hb]), (rank, spread, call)
        state["step"] += 1
dist.barrier()
dist.destroy_process_group()
print("ok", rank)
'''


@pytest.mark.parametrize("world", [2, 3])
def test_halo_exchange_gloo(tmp_path, world):
    script = tmp_path / "worker.py"
    script.write_text(GLOO_WORKER)
    port = 29500 + world + (os.getpid() % 500)
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}",
                        "--master-addr", "127.0.0.1", "--master-port", str(port), str(script), ROOT],
                       capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-3000:]
    assert r.stdout.count("ok") == world


def test_host_pipeline_band_schedule(pkg, monkeypatch):
    """The row-band schedule of chaqsdf_u8 (host logic, no device): bands cover the image exactly, are multiples of
    32 rows except the last, every band is tall enough to finish some output rows (>= spread + 32), short bands
    come first and last, and small images or spreads beyond the banded path are not pipelined."""
    for var in ("CHAQSDF_PIPE_ROWS", "CHAQSDF_PIPE_FIRST"):
        monkeypatch.delenv(var, raising=False)
    pkg.reload_tuning()
    assert pkg.pipeline_plan(1024, 1024, 1, 64) == []          # < 4 Mpx
    assert pkg.pipeline_plan(16384, 16384, 1, 1024) == []      # general path (spread > 256)
    for (w, h, stride, spread) in [(16384, 16384, 1, 256), (16384, 16384, 2, 256), (4096, 4096, 1, 64), (2048, 2048, 1, 40),
                                   (32768, 4000, 1, 8), (3000, 5000, 4, 200), (2048, 2049, 1, 256)]:
        bands = pkg.pipeline_plan(w, h, stride, spread)
        if not bands:
            continue
        assert sum(bands) == h, (w, h, bands)
        assert all(b % 32 == 0 for b in bands[:-1]) and all(b > 0 for b in bands), bands
        assert all(b >= spread + 32 for b in bands[:-1]), (spread, bands)
        assert bands[0] <= max(bands) and len(bands) >= 2
        assert max(bands) * w * stride <= (48 << 20) or max(bands) <= 2 * spread + 96, "a band is ~16 MB of input"
    bands = pkg.pipeline_plan(16384, 16384, 1, 256)
    assert len(bands) >= 8 and bands[0] < bands[len(bands) // 2] and bands[-1] < bands[len(bands) // 2], bands
    monkeypatch.setenv("CHAQSDF_PIPE_ROWS", "160")
    pkg.reload_tuning()
    b = pkg.pipeline_plan(2048, 2048, 1, 40)
    assert sum(b) == 2048 and max(b) <= 160, b                # 2*spread + 64 = 144 <= 160
    monkeypatch.setenv("CHAQSDF_PIPE_ROWS", "64")             # too short for the context: raised to 2*spread + 64
    pkg.reload_tuning()
    b = pkg.pipeline_plan(2048, 2048, 1, 40)
    assert sum(b) == 2048 and all(x >= 40 + 32 for x in b[:-1]), b
    monkeypatch.delenv("CHAQSDF_PIPE_ROWS")
    pkg.reload_tuning()


def test_reference_main_links_against_the_dropin(pkg, tmp_path):
    """INTEGRATION.md section 1 as a build: the reference's own program (openmp/sdfgen.c, unmodified, where it lies in
    the reference tree) compiled WITHOUT openmp/df.c and linked against libchaqsdf_b200.so; dist_transform_2d must be
    the only thing it takes from the library. (Running it is a GPU test: test_reference_main_runs_on_the_dropin.)"""
    ref_c = "/root/reference/openmp/sdfgen.c"
    if not os.path.exists(ref_c):
        pytest.skip("reference tree absent")
    exe = tmp_path / "chaq_sdfgen"
    libdir = os.path.join(ROOT, "chaq-sdfgen_b200")
    r = subprocess.run(["/usr/bin/gcc", "-std=c11", "-O3", "-fopenmp", "-w", "-I/root/reference/include", ref_c, "-L" + libdir,
                        "-lchaqsdf_b200", "-Wl,-rpath," + libdir, "-lm", "-o", str(exe)], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr[-2000:]
    und = subprocess.run(["nm", "-D", "--undefined-only", str(exe)], capture_output=True, text=True).stdout
    assert "dist_transform_2d" in und and "chaqsdf_" not in und
    r = subprocess.run([str(exe), "-h"], capture_output=True, text=True, timeout=60)  # the reference's own usage text
    assert r.returncode == 0 and r.stdout.startswith("usage: chaq_sdfgen")


def test_syn_glyph_generators_agree(oracle):
    """SYN-GLYPH exists twice (C in the oracle for the tests, numpy in bench.py for the GPU box): same bytes."""
    import bench
    for n, first, count in ((256, 0, 12), (256, 4090, 6), (128, 3, 5), (96, 0, 4)):
        a, b = oracle.syn_glyphs(count, n, first), bench.syn_glyphs(count, n, first)
        assert np.array_equal(a, b), (n, first)
        assert not a[:, :16].any() and not a[:, -16:].any() and not a[:, :, :16].any() and not a[:, :, -16:].any()
    cov = (oracle.syn_glyphs(64) > 0).mean()
    assert 0.05 < cov < 0.5


def test_peer_halo_pull_plan_is_paired_and_complete(pkg):
    """The peer-memory exchange (pull_halos) on paper: every signal has exactly one waiter, only owned rows are read,
    and after the copies every rank's buffer holds image rows [r0 - halo_top, r1 + halo_bot) in order -- including
    spreads larger than a band (multi-hop halos) and ragged band heights."""
    for (h, world, spread) in [(64, 2, 5), (100, 3, 40), (37, 4, 3), (200, 8, 100), (33, 5, 64), (16, 4, 1), (50, 7, 50)]:
        plans = [pkg.halo_plan(h, world, r, spread) for r in range(world)]
        lists = [pkg.halo_pull_list(plans, r) for r in range(world)]
        bufs = []
        for (r0, r1, ht, hb, _, _) in plans:
            b = np.full(ht + (r1 - r0) + hb, -1, np.int64)
            b[ht:ht + (r1 - r0)] = np.arange(r0, r1)  # pass 1 writes the owned rows
            bufs.append(b)
        for r in range(world):
            signal_to, wait_for, copies = lists[r]
            for p in signal_to:
                assert r in lists[p][1], (h, world, spread, r, p)
            for p in wait_for:
                assert r in lists[p][0], (h, world, spread, r, p)
            assert sorted({c[0] for c in copies}) == wait_for
        snapshot = [b.copy() for b in bufs]  # pulls read owned rows only, so their order cannot matter
        for r in range(world):
            for (peer, d0, d1, s0, s1) in lists[r][2]:
                o0, o1, oht = plans[peer][0], plans[peer][1], plans[peer][2]
                assert oht <= s0 < s1 <= oht + (o1 - o0), "reads rows the peer does not own"
                bufs[r][d0:d1] = snapshot[peer][s0:s1]
        for r, (r0, r1, ht, hb, _, _) in enumerate(plans):
            assert np.array_equal(bufs[r], np.arange(r0 - ht, r1 + hb)), (h, world, spread, r)


# ---------------------------------------------------------------- build / merge / paint (hull_core.h, paint_core.h), host emulation
def _build_emul3(tmp_path_factory, name, *defines):
    so = str(tmp_path_factory.mktemp(name) / f"lib{name}.so")
    subprocess.run(["/usr/bin/g++", "-O2", "-std=c++17", "-ffp-contract=off", "-fPIC", "-shared", *defines, "-o", so,
                    os.path.join(ROOT, "tests", "emul", "emul3.cpp")], check=True)
    L = C.CDLL(so)
    u8p = np.ctypeslib.ndpointer(np.uint8, flags="C")
    L.emul3_sdf.argtypes = [u8p, C.c_int, C.c_int, C.c_size_t, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, u8p,
                            C.POINTER(C.c_int), C.c_int, C.c_int]
    L.emul3_sdf_wide.argtypes = L.emul3_sdf.argtypes
    return L


@pytest.fixture(scope="session")
def emul3(tmp_path_factory):
    return _build_emul3(tmp_path_factory, "emul3")


class _Wide:
    """The wide instantiation of the same library (16-bit row distances, 64-bit domination test, L <= 1024)."""

    def __init__(self, L):
        self.emul3_sdf = L.emul3_sdf_wide


@pytest.fixture(scope="session")
def emul3_wide(emul3):
    return _Wide(emul3)


def _emul3_u8(L, mask, spread, asym, build_blocks, emit_rows, y0=0, y1=None, order=None, nsub=1):
    """order None: both junction orders (top to bottom; in rounds) and both build shapes (whole runs; four sub-runs
    repaired inside the run first, as a CTA of the build kernel does) must give the same bytes."""
    if order is None:
        a, ce = _emul3_u8(L, mask, spread, asym, build_blocks, emit_rows, y0, y1, 0)
        b, _ = _emul3_u8(L, mask, spread, asym, build_blocks, emit_rows, y0, y1, 1)
        c, _ = _emul3_u8(L, mask, spread, asym, build_blocks, emit_rows, y0, y1, 1, 4)
        d, _ = _emul3_u8(L, mask, spread, asym, build_blocks, emit_rows, y0, y1, 0, 3)
        assert np.array_equal(a, b), "junction orders disagree"
        assert np.array_equal(a, c) and np.array_equal(a, d), "sub-run builds disagree"
        return a, ce
    h, w = mask.shape
    y1 = h if y1 is None else y1
    out = np.zeros((h, w), np.uint8)
    ce = C.c_int()
    rc = L.emul3_sdf(np.ascontiguousarray(mask), w, h, spread, int(asym), build_blocks, emit_rows, y0, y1, out, C.byref(ce),
                     order, nsub)
    assert rc == 0, f"emulated build/merge/emit failed (rc={rc}: 1 = cap_eff beyond the build's limit, 2 = row count, >=1000 = out-of-range access)"
    return out, ce.value


def test_hull_build_merge_paint_matches_oracle_random(emul3, oracle):
    """Run lengths from one block (a junction every 32 rows, far shorter than the reach) to 32 blocks."""
    rng = np.random.default_rng(5)
    for it in range(100):
        h, w = int(rng.integers(1, 400)), int(rng.integers(1, 24))
        dens = rng.choice([0.0, 0.01, 0.05, 0.3, 0.5, 0.9, 1.0])
        mask = (rng.random((h, w)) < dens).astype(np.uint8)
        if it % 3 == 0:
            yy, xx = np.mgrid[0:h, 0:w]
            mask = np.zeros((h, w), np.uint8)
            for _ in range(3):
                cy, cx, r = rng.integers(0, h), rng.integers(0, w), rng.integers(1, 60)
                mask |= ((yy - cy) ** 2 + (xx - cx) ** 2 <= r * r).astype(np.uint8)
        img = (mask * 255).astype(np.uint8)
        for spread in (1, 2, 5, 16, 40, 100, 256):
            asym = bool(it & 1)
            want = oracle.sdf(img, w, h, spread=spread, asymmetric=asym)
            for bb, er in ((1, 0), (2, 7), (3, 33), (8, 64), (32, 100)):
                got, _ = _emul3_u8(emul3, mask, spread, asym, bb, er)
                assert np.array_equal(got, want), (it, h, w, spread, bb, er)


def test_hull_discs_noise_and_row_ranges(emul3, oracle):
    for n, spread, bb, er in ((384, 16, 2, 0), (384, 64, 4, 100), (1024, 256, 16, 400), (1024, 256, 1, 37), (700, 100, 3, 64), (1024, 33, 3, 50), (1024, 15, 1, 0)):
        img = oracle.syn_discs(n)
        got, ce = _emul3_u8(emul3, (img > 127).astype(np.uint8), spread, False, bb, er)
        assert ce <= min(spread + 1, 256)
        assert np.array_equal(got, oracle.sdf(img, n, n, spread=spread)), (n, spread, bb, er)
    noise = oracle.syn_noise(300)
    for spread in (3, 8, 50, 256):
        got, _ = _emul3_u8(emul3, (noise > 127).astype(np.uint8), spread, True, 2, 50)
        assert np.array_equal(got, oracle.sdf(noise, 300, 300, spread=spread, asymmetric=True)), spread
    # a row band of a larger buffer (unaligned ends), as the multi-GPU and pipelined callers use it
    img = oracle.syn_discs(512)
    want = oracle.sdf(img, 512, 512, spread=40)
    got, _ = _emul3_u8(emul3, (img > 127).astype(np.uint8), 40, False, 4, 96, 100, 333)
    assert np.array_equal(got[100:333], want[100:333])


def test_hull_junction_repair_removes_long_chains(emul3, oracle):
    """A junction whose repair removes ~200 sites on one side: above row J every row of a column is a site of height
    (200 - x)^2 (a vertical edge 200 columns away), from row J on the edge is adjacent (height 1), which dominates them
    all. The repair finds how far to go by galloping + bisection inside the adjacent build run and site by site beyond
    it; run lengths of 16, 8 and 4 blocks put the end of the chain inside the run, one run further and several runs
    further. Mirrored vertically for the other side of the junction."""
    for J, h in ((512, 1024), (256, 700), (384, 640)):
        mask = np.zeros((h, 230), np.uint8)
        mask[:J, 200:] = 1
        mask[J:, 1:] = 1
        for m in (mask, np.ascontiguousarray(mask[::-1])):
            img = (m * 255).astype(np.uint8)
            want = oracle.sdf(img, 230, h, spread=256)
            for bb in (16, 8, 4):
                got, _ = _emul3_u8(emul3, m, 256, False, bb, 100)
                assert np.array_equal(got, want), (J, h, bb)


def test_hull_merge_rounds_never_share_a_hull_word(emul3, oracle):
    """Junctions repaired in the same round run concurrently on the GPU with plain loads and stores, so two of them
    must never store the same 32-row hull word. Counting rounds in rows (ceil(4L / run length)) allowed exactly that
    for 2L mod 32 in 2..16 with short runs (spread 17..24 with runs of 3 blocks, spread ~100 with 13); rounds are
    counted in words now (hull_merge_phases). The emulation fails with rc 2500 when two same-round junctions store one
    word; dense glyph-like data makes the repairs reach as far as they can."""
    rng = np.random.default_rng(21)
    h, w = 1500, 6
    masks = [(rng.random((h, w)) < 0.5).astype(np.uint8), (rng.random((h, w)) < 0.03).astype(np.uint8)]
    yy, xx = np.mgrid[0:h, 0:w]
    m = np.zeros((h, w), np.uint8)
    for _ in range(40):
        cy, cx, r = rng.integers(0, h), rng.integers(-30, w + 30), rng.integers(2, 40)
        m |= ((yy - cy) ** 2 + (xx - cx) ** 2 <= r * r).astype(np.uint8)
    masks.append(m)
    for mask in masks:
        img = (mask * 255).astype(np.uint8)
        for spread, bbs in ((17, (3, 1, 2)), (20, (3, 5)), (24, (3, 4)), (100, (13, 7)), (256, (16, 17, 31)), (1, (1,)), (8, (1, 2))):
            want = oracle.sdf(img, w, h, spread=spread)
            for bb in bbs:
                got, _ = _emul3_u8(emul3, mask, spread, False, bb, 0, order=1)
                assert np.array_equal(got, want), (spread, bb)


def test_hull_wide_variant_matches_oracle_beyond_the_banded_spreads(emul3_wide, oracle):
    """Groundwork for moving spreads > 256 (config 5) from the first-generation sweep onto build / merge / emit: the
    same core compiled wide must reproduce the oracle for reaches up to 1024 rows, for run lengths from far below
    the reach to 128 blocks, and must agree with the narrow build where both apply."""
    rng = np.random.default_rng(9)
    for it in range(12):
        h, w = int(rng.integers(200, 2600)), int(rng.integers(1, 6))
        if it % 2 == 0:
            yy, xx = np.mgrid[0:h, 0:w]
            mask = np.zeros((h, w), np.uint8)
            for _ in range(4):
                cy, cx, r = rng.integers(0, h), rng.integers(-40, w + 40), rng.integers(1, 500)
                mask |= ((yy - cy) ** 2 + (xx - cx) ** 2 <= r * r).astype(np.uint8)
        else:
            mask = (rng.random((h, w)) < rng.choice([0.0005, 0.003, 0.02, 0.5])).astype(np.uint8)
        img = (mask * 255).astype(np.uint8)
        for spread in (300, 700, 1024):
            asym = bool(it & 1)
            want = oracle.sdf(img, w, h, spread=spread, asymmetric=asym)
            for bb, er in ((4, 0), (16, 333), (64, 1000), (128, 0)):
                got, ce = _emul3_u8(emul3_wide, mask, spread, asym, bb, er)
                assert ce <= spread + 1
                assert np.array_equal(got, want), (it, h, w, spread, bb, er)
    img = oracle.syn_discs(640)
    m = (img > 127).astype(np.uint8)
    for spread in (100, 256, 600):
        got, _ = _emul3_u8(emul3_wide, m, spread, False, 8, 200)
        assert np.array_equal(got, oracle.sdf(img, 640, 640, spread=spread)), spread


# ---------------------------------------------------------------- chaq_sdfgen command line (reference sdfgen.c:122-244)
CLI = os.path.join(ROOT, "chaq-sdfgen_b200", "bin", "chaq_sdfgen")


def _cli(*args, stdin=None):
    return subprocess.run([CLI, *args], input=stdin, capture_output=True, timeout=120)


def test_cli_argument_contract(pkg):
    if not os.path.exists(CLI):
        pytest.skip("bin/chaq_sdfgen not built (needs the stb headers of the reference tree)")
    r = _cli("-h")
    assert r.returncode == 0 and r.stdout.startswith(b"usage: chaq_sdfgen [-f filetype] -i file -o file [-q n] [-s n] [-ahln]")
    r = _cli("-alh")  # flags bundle; h anywhere in it prints the usage and exits 0 (sdfgen.c:206-209)
    assert r.returncode == 0 and b"usage:" in r.stdout
    for args, msg in ((("-i", "a.png", "-o", "b.png", "-s", "0"), b"Invalid value given for spread"),
                      (("-i", "a.png", "-o", "b.png", "-q", "0"), b"Invalid value given for jpeg quality"),
                      (("-i", "a.png", "-o", "b.png", "-q", "101"), b"Invalid value given for jpeg quality"),
                      (("-o", "b.png"), b"No input file specified."),
                      (("-i", "a.png"), b"No output file specified."),
                      (("-i",), b"No input file specified."),
                      (("-i", "a.png", "-o", "b.png", "-f", "gif"), b"Invalid filetype specified."),
                      (("-i", "a.png", "-o", "b.png", "-s"), b"No number specified with spread.")):
        r = _cli(*args)
        assert r.returncode == 255, args          # exit(-1), sdfgen.c:24-30
        assert msg in r.stderr and b"usage:" in r.stdout, args
    r = _cli("-i", "/nonexistent/x.png", "-o", "/tmp/y.png")
    assert r.returncode == 255 and b"Input file could not be opened." in r.stderr and b"usage:" not in r.stdout


def test_atlas_pack_host_only(pkg):
    """chaqsdf_atlas_pack: rectangles inside the atlas, pairwise disjoint including their padding,
    deterministic, reasonably dense; degenerate inputs."""
    rng = np.random.default_rng(5)
    for (n, aw, pad, hi) in ((500, 1024, 2, 200), (4096, 4096, 1, 120), (50, 64, 0, 64), (1, 10, 3, 10)):
        wh = rng.integers(1, hi + 1, (n, 2)).astype(np.uint32)
        wh[:, 0] = np.minimum(wh[:, 0], aw)
        if n > 10:
            wh[3] = (0, 7)      # empty rectangles are allowed
            wh[7] = (aw, 1)     # as wide as the atlas
        xy, ah = pkg.atlas_pack(wh, aw, pad)
        xy2, ah2 = pkg.atlas_pack(wh, aw, pad)
        assert ah == ah2 and np.array_equal(xy, xy2)
        occ = np.zeros((ah + pad + 1, aw + pad + 1), np.int32)
        area = 0
        for (w_, h_), (x, y) in zip(wh.tolist(), xy.tolist()):
            if w_ == 0 or h_ == 0:
                assert (x, y) == (0, 0)
                continue
            assert x + w_ <= aw and y + h_ <= ah
            occ[y:y + h_ + pad, x:x + w_ + pad] += 1   # the rectangle and its padding
            area += w_ * h_
        assert (occ[:, :aw] <= 1).all(), "rectangles (with padding) overlap"
        assert ah == max(y + h_ for (w_, h_), (x, y) in zip(wh.tolist(), xy.tolist()) if w_ and h_)
        if n >= 500:
            assert area / (aw * ah) > 0.75, area / (aw * ah)
    with pytest.raises(pkg.ChaqSdfError):
        pkg.atlas_pack(np.array([[11, 1]], np.uint32), 10, 0)
