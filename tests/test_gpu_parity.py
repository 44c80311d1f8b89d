"""GPU parity (run on the B200 box: pytest -m gpu). Every call goes through the C ABI (libchaqsdf_b200.so);
the oracle (oracle/liboracle.so) and the prebuilt stock reference (oracle/_ref/libchaqref.so) are the checkers.
Bar: bit-exact bytes and squared distances; float fields 0 ulp where D2 < 2^24 (every case below)."""
import os
import sys

import numpy as np
import pytest

from conftest import GOLDEN, ROOT

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


def gpu_u8(pkg, img, w, h, **kw):
    d = torch.from_numpy(np.ascontiguousarray(img)).cuda()
    out = pkg.sdf_u8_device(d, w, h, **kw)
    torch.cuda.synchronize()
    return out.cpu().numpy()


def random_shapes(rng, h, w):
    yy, xx = np.mgrid[0:h, 0:w]
    mask = np.zeros((h, w), bool)
    for _ in range(int(rng.integers(1, 8))):
        cy, cx, r = rng.integers(0, h), rng.integers(0, w), rng.integers(1, max(2, min(h, w) // 2))
        m = (yy - cy) ** 2 + (xx - cx) ** 2 <= r * r
        mask = (mask & ~m) if rng.random() < 0.3 else (mask | m)
    return (mask * 255).astype(np.uint8)


def test_golden_pair_host_api(pkg):
    z = np.load(os.path.join(GOLDEN, "sample_pair.npz"))
    out = pkg.sdf_u8(np.ascontiguousarray(z["ya"]), 200, 200, stride=2, offset=0, test_above=True, spread=100,
                     asymmetric=True)
    assert np.array_equal(out, z["out"])


def test_ref_vectors(pkg):
    z = np.load(os.path.join(GOLDEN, "ref_vectors.npz"))
    g = np.ascontiguousarray(z["grad_ya"])
    n = g.shape[0]
    flags = {"default": (1, True, False), "l": (0, True, False), "n": (1, False, False), "a": (1, True, True),
             "ln": (0, False, False), "aln": (0, False, True)}
    for name, (off, above, asym) in flags.items():
        for s in (7, 64):
            got = gpu_u8(pkg, g, n, n, stride=2, offset=off, test_above=above, spread=s, asymmetric=asym)
            assert np.array_equal(got, z[f"grad_{name}_s{s}"]), (name, s)
            assert np.array_equal(pkg.sdf_u8(g, n, n, stride=2, offset=off, test_above=above, spread=s,
                                             asymmetric=asym), z[f"grad_{name}_s{s}"])


def test_small_shapes_and_edges(pkg, oracle):
    rng = np.random.default_rng(21)
    shapes = [(1, 1), (1, 2), (2, 1), (2, 2), (1, 33), (33, 1), (7, 513), (513, 7), (31, 31), (32, 32), (33, 33),
              (100, 1025), (257, 300), (300, 200), (200, 300), (64, 2048)]
    for (h, w) in shapes:
        for kind in range(4):
            if kind == 0:
                img = (rng.random((h, w)) < 0.5).astype(np.uint8) * 255
            elif kind == 1:
                img = (rng.random((h, w)) < 0.02).astype(np.uint8) * 255
            elif kind == 2:
                img = random_shapes(rng, h, w)
            else:
                img = rng.integers(0, 256, (h, w), dtype=np.uint8)  # raw bytes: exercises the 127/128 threshold
            for spread in (1, 3, 16, 64, 300):
                for asym in (False, True):
                    above = bool((kind + spread) & 1) or kind != 3
                    got = gpu_u8(pkg, img, w, h, spread=spread, asymmetric=asym, test_above=above)
                    want = oracle.sdf(img, w, h, spread=spread, asymmetric=asym, test_above=above)
                    assert np.array_equal(got, want), (h, w, kind, spread, asym, int((got != want).sum()))


def test_degenerates(pkg, oracle):
    n = 64
    for val in (0, 126, 127, 128, 255):
        img = np.full((n, n), val, np.uint8)
        for above in (True, False):
            got = gpu_u8(pkg, img, n, n, spread=8, test_above=above)
            assert np.array_equal(got, oracle.sdf(img, n, n, spread=8, test_above=above)), (val, above)
    for corner in ((0, 0), (0, n - 1), (n - 1, 0), (n - 1, n - 1)):
        img = np.zeros((n, n), np.uint8)
        img[corner] = 255
        for spread in (4, 100):
            assert np.array_equal(gpu_u8(pkg, img, n, n, spread=spread), oracle.sdf(img, n, n, spread=spread))
            assert np.array_equal(gpu_u8(pkg, 255 - img, n, n, spread=spread), oracle.sdf(255 - img, n, n, spread=spread))
    chk = ((np.indices((n, n)).sum(0) & 1) * 255).astype(np.uint8)
    assert np.array_equal(gpu_u8(pkg, chk, n, n, spread=5), oracle.sdf(chk, n, n, spread=5))
    half = np.zeros((n, n), np.uint8)
    half[:, n // 2:] = 255
    assert np.array_equal(gpu_u8(pkg, half, n, n, spread=20), oracle.sdf(half, n, n, spread=20))
    assert np.array_equal(gpu_u8(pkg, half.T.copy(), n, n, spread=20), oracle.sdf(half.T.copy(), n, n, spread=20))


def test_band_sizes_do_not_change_the_result(pkg, oracle):
    img = oracle.syn_discs(1024)
    want = oracle.sdf(img, 1024, 1024, spread=48)
    try:
        for band in ("1", "17", "64", "333", "5000"):
            os.environ["CHAQSDF_BAND_ROWS"] = band
            pkg.reload_tuning()
            assert np.array_equal(gpu_u8(pkg, img, 1024, 1024, spread=48), want), band
    finally:
        os.environ.pop("CHAQSDF_BAND_ROWS", None)
        pkg.reload_tuning()


def test_strided_input_and_invert(pkg, oracle):
    rng = np.random.default_rng(4)
    h, w = 130, 1100
    ya = rng.integers(0, 256, (h, w, 2), dtype=np.uint8)
    ya[20:90, 100:900, :] = 200
    for off in (0, 1):
        for above in (True, False):
            got = gpu_u8(pkg, ya, w, h, stride=2, offset=off, test_above=above, spread=12)
            assert np.array_equal(got, oracle.sdf(ya, w, h, stride=2, offset=off, test_above=above, spread=12))
    rgba = rng.integers(0, 256, (h, w, 4), dtype=np.uint8)
    got = gpu_u8(pkg, rgba, w, h, stride=4, offset=3, spread=9)
    assert np.array_equal(got, oracle.sdf(rgba, w, h, stride=4, offset=3, spread=9))


def test_exact_fields(pkg, oracle):
    rng = np.random.default_rng(8)
    for (h, w) in ((64, 64), (200, 333), (512, 512), (1, 50), (50, 1)):
        for img in (random_shapes(rng, h, w), (rng.random((h, w)) < 0.01).astype(np.uint8) * 255,
                    np.zeros((h, w), np.uint8), np.full((h, w), 255, np.uint8)):
            d = torch.from_numpy(img).cuda()
            res = pkg.sdf_fields_device(d, w, h, spread=16)
            torch.cuda.synchronize()
            out, sdf, din, dout = oracle.sdf(img, w, h, spread=16, fields=True)
            assert np.array_equal(res["d2_in"].cpu().numpy(), din)
            assert np.array_equal(res["d2_out"].cpu().numpy(), dout)
            assert np.array_equal(res["out"].cpu().numpy(), out)
            assert np.array_equal(res["sdf"].cpu().numpy().view(np.uint32), sdf.view(np.uint32))  # 0 ulp
            fin = np.where(din == 2**31 - 1, np.float32(np.inf), np.sqrt(din.astype(np.float32)))
            assert np.array_equal(res["f_in"].cpu().numpy().view(np.uint32), fin.astype(np.float32).view(np.uint32))


def test_dist_transform_2d_dropin(pkg, oracle, ref):
    rng = np.random.default_rng(9)
    for n in (64, 500, 1024):
        seeds = rng.random((n, n)) < 0.002
        f = np.where(seeds, np.float32(0), np.float32(np.inf)).astype(np.float32)
        got = pkg.dist_transform_2d(f.copy())
        assert np.array_equal(got.view(np.uint32), ref.dist_transform_2d(f).view(np.uint32))  # stock build, 0 ulp
    f = np.full((40, 70), np.inf, np.float32)  # non-square (stock build is UB here), no seed at all
    assert np.isinf(pkg.dist_transform_2d(f.copy())).all()
    f[3, 60] = 0
    got = pkg.dist_transform_2d(f.copy())
    assert np.array_equal(got.view(np.uint32), oracle.dist_transform_2d_binary(f).view(np.uint32))


def test_dist_transform_2d_general_heights(pkg, ref, monkeypatch):
    """Sampled functions (any finite height or +INF) take the float32 path that restates df.c:29-97 operation by
    operation: bit-identical to the stock OpenMP build on square images (the only shape it is defined for)."""
    rng = np.random.default_rng(31)
    for n in (1, 2, 3, 17, 64, 257, 700):
        for kind in range(4):
            f = (rng.random((n, n)) * [1.0, 50.0, 4000.0, 1e6][kind]).astype(np.float32)
            if kind >= 1:
                f[rng.random((n, n)) < [0, 0.3, 0.9, 0.995][kind]] = np.inf
            if kind == 2:
                f[rng.random((n, n)) < 0.05] = -3.5  # negative heights: accepted like the stock build accepts them
            want = ref.dist_transform_2d(f)
            got = pkg.dist_transform_2d(f.copy())
            nan = np.isnan(want)  # sqrtf of a negative envelope value: NaN in both, payload bits are not compared
            assert np.array_equal(np.isnan(got), nan), (n, kind)
            assert np.array_equal(got.view(np.uint32)[~nan], want.view(np.uint32)[~nan]), (n, kind)
    f = np.full((33, 33), np.inf, np.float32)
    f[5, 7] = 2.25
    assert np.array_equal(pkg.dist_transform_2d(f.copy()).view(np.uint32), ref.dist_transform_2d(f).view(np.uint32))
    # non-square: the stock build is undefined there; check against a float64 brute force
    h, w = 23, 41
    f = (rng.random((h, w)) * 30).astype(np.float32)
    f[rng.random((h, w)) < 0.7] = np.inf
    yy, xx = np.mgrid[0:h, 0:w]
    fin = np.isfinite(f)
    cand = ((yy[..., None] - yy[fin]) ** 2 + (xx[..., None] - xx[fin]) ** 2 + f[fin].astype(np.float64))
    want = np.sqrt(cand.min(axis=-1))
    got = pkg.dist_transform_2d(f.copy())
    assert np.allclose(got, want, rtol=1e-5, atol=1e-5)
    # stock-compatible mode on a binary image beyond side 4098, where the stock float32 envelope is inexact
    # (SURVEY.md 0.3): same bits as the stock build; the default (exact) mode differs from it there
    n = 4300
    seeds = rng.random((n, n)) < 0.00002
    seeds[0, 0] = True
    f = np.where(seeds, np.float32(0), np.float32(np.inf)).astype(np.float32)
    want = ref.dist_transform_2d(f)
    monkeypatch.setenv("CHAQSDF_DT2D", "stock")
    pkg.reload_tuning()
    got = pkg.dist_transform_2d(f.copy())
    monkeypatch.delenv("CHAQSDF_DT2D")
    pkg.reload_tuning()
    assert np.array_equal(got.view(np.uint32), want.view(np.uint32))


def test_batched_glyphs(pkg, oracle):
    rng = np.random.default_rng(12)
    n, s = 37, 64
    imgs = np.stack([random_shapes(rng, s, s) for _ in range(n)])
    d = torch.from_numpy(imgs).cuda()
    out = pkg.sdf_u8_batched_device(d, n, s, s, spread=16)
    torch.cuda.synchronize()
    got = out.cpu().numpy()
    for k in range(n):
        assert np.array_equal(got[k], oracle.sdf(imgs[k], s, s, spread=16)), k


def _batched(pkg, imgs, **kw):
    n, h, w = imgs.shape
    d = torch.from_numpy(np.ascontiguousarray(imgs)).cuda()
    out = pkg.sdf_u8_batched_device(d, n, w, h, **kw)
    torch.cuda.synchronize()
    return out.cpu().numpy()


def test_fused_small_image_kernel(pkg, oracle, monkeypatch):
    """Batches of small images with spread <= 32 go through glyph_kernel (one CTA per image, both passes in shared
    memory). Every window size of the kernel (L <= 8 / 16 / 32), both threshold directions, asymmetric, ragged heights,
    every width it takes, and a single image forced through it; each image against the oracle, and the general
    two-pass path (CHAQSDF_GLYPH=never) must agree byte for byte."""
    rng = np.random.default_rng(31)
    before = pkg.launch_count()
    cases = [(5, 32, 32, 4, {}), (9, 64, 100, 8, {"asymmetric": True}), (7, 256, 33, 16, {}), (6, 128, 128, 32, {}),
             (4, 96, 70, 16, {"test_above": False}), (3, 256, 256, 32, {"asymmetric": True}), (2, 160, 1, 3, {}),
             (8, 224, 75, 1, {}), (3, 256, 300, 20, {})]
    for (n, w, h, spread, kw) in cases:
        imgs = np.stack([random_shapes(rng, h, w) if k % 3 else (rng.random((h, w)) < 0.5).astype(np.uint8) * 255
                         for k in range(n)])
        if n > 2:
            imgs[0][:] = 0
        if n > 3:
            imgs[1][:] = 255
            imgs[2][h // 2, w // 2] ^= 255
        got = _batched(pkg, imgs, spread=spread, **kw)
        for k in range(n):
            assert np.array_equal(got[k], oracle.sdf(imgs[k], w, h, spread=spread, **kw)), (n, w, h, spread, kw, k)
        monkeypatch.setenv("CHAQSDF_GLYPH", "never")
        pkg.reload_tuning()
        two_pass = _batched(pkg, imgs, spread=spread, **kw)
        monkeypatch.delenv("CHAQSDF_GLYPH")
        pkg.reload_tuning()
        assert np.array_equal(got, two_pass), (n, w, h, spread)
    # the fused kernel is one launch per batch: 9 batches fused + 9 x (pass 1 + build + paint [+ merges]) unfused
    assert pkg.launch_count() - before >= 9 + 9 * 3
    # one image, forced
    img = random_shapes(rng, 256, 256)
    monkeypatch.setenv("CHAQSDF_GLYPH", "always")
    pkg.reload_tuning()
    l0 = pkg.launch_count()
    got = gpu_u8(pkg, img, 256, 256, spread=16)
    assert pkg.launch_count() - l0 == 1
    monkeypatch.delenv("CHAQSDF_GLYPH")
    pkg.reload_tuning()
    assert np.array_equal(got, oracle.sdf(img, 256, 256, spread=16))


def test_config3_at_its_shape(pkg, oracle):
    """BASELINE config 3: 4096 glyphs of 256 x 256 (SYN-GLYPH of SURVEY.md 8(d), oracle/edt_oracle.c), spread 16, one
    batched call; every glyph against the oracle."""
    n, s, spread = 4096, 256, 16
    imgs = oracle.syn_glyphs(n, s)
    got = _batched(pkg, imgs, spread=spread)
    bad = [k for k in range(n) if not np.array_equal(got[k], oracle.sdf(imgs[k], s, s, spread=spread))]
    assert not bad, f"{len(bad)} glyphs differ, first {bad[:5]}"


def test_row_band_sharding_single_process(pkg, oracle):
    """The multi-GPU decomposition, emulated on one GPU: each 'rank' runs pass 1 on its band, halos are copied
    as exchange_halos would, pass 2 runs per band; the stitched result must equal the unsharded one."""
    for (h, w, spread) in ((700, 640, 40), (900, 333, 300)):  # banded and general intermediate
      img = oracle.syn_discs(w, h)
      want = oracle.sdf(img, w, h, spread=spread)
      d_img = torch.from_numpy(img).cuda()
      for world in (2, 3, 8, 32):
          inters = {}
          for rank in range(world):
              r0, r1 = pkg.shard_rows(h, world, rank)
              buf = torch.empty((r1 - r0, pkg.inter_row_bytes(w, spread)), dtype=torch.uint8, device="cuda")
              pkg.rowscan_device(d_img[r0:r1], w, r1 - r0, spread, buf)
              inters[rank] = buf
          full = torch.cat([inters[r] for r in range(world)])
          for rank in range(world):
              r0, r1, ht, hb, recvs, sends = pkg.halo_plan(h, world, rank, spread)
              buf = full[r0 - ht:r1 + hb].contiguous()
              out = torch.empty((r1 - r0, w), dtype=torch.uint8, device="cuda")
              pkg.envelope_device(buf, w, buf.shape[0], ht, ht + r1 - r0, spread, False, out)
              torch.cuda.synchronize()
              assert np.array_equal(out.cpu().numpy(), want[r0:r1]), (world, rank)


def test_transpose_u8(pkg):
    """chaqsdf_transpose_u8_device on aligned, ragged and unaligned operands, against torch."""
    g = torch.Generator(device="cuda").manual_seed(3)
    for (rows, cols, off, sp_extra, dp_extra) in ((256, 128, 0, 0, 0), (64, 64, 0, 0, 0), (257, 333, 0, 0, 0), (1000, 70, 5, 3, 1),
                                                  (33, 2049, 16, 16, 0), (1, 1, 0, 0, 0)):
        src_pitch, dst_pitch = cols + off + sp_extra, rows + dp_extra
        src = torch.randint(0, 256, (rows, src_pitch), dtype=torch.uint8, device="cuda", generator=g)
        dst = torch.full((cols, dst_pitch), 0xAB, dtype=torch.uint8, device="cuda")
        pkg.transpose_u8_device(src.data_ptr() + off, src_pitch, rows, cols, dst.data_ptr(), dst_pitch)
        torch.cuda.synchronize()
        assert torch.equal(dst[:, :rows], src[:, off:off + cols].t()), (rows, cols, off)
        assert bool((dst[:, rows:] == 0xAB).all()), "wrote past the destination rows"


def test_transposed_band_sharding_single_process(pkg, oracle):
    """The transposed-band decomposition (sdf_u8_sharded_t), emulated on one GPU: each 'rank' gets its band plus
    `spread` rows of the INPUT on either side, as the peer pull delivers them, and works alone; its rows must equal
    the unsharded result. Covers bands thinner than the spread (context from several neighbours), ragged band heights
    (column ranges that do not start at a multiple of 32) and the wide (16-bit) intermediate."""
    for (h, w, spread) in ((700, 640, 40), (515, 333, 100), (1100, 200, 300)):
        img = oracle.syn_discs(w, h)
        want = oracle.sdf(img, w, h, spread=spread)
        d_img = torch.from_numpy(img).cuda()
        for world in (2, 3, 8, 32):
            bufs = {}
            for rank in range(world):
                r0, r1, ht, hb, _, _ = pkg.halo_plan(h, world, rank, spread)
                ext = d_img[r0 - ht:r1 + hb].contiguous()
                out = pkg.sdf_u8_tband_device(ext, w, ext.shape[0], ht, r1 - r0, spread, bufs=bufs)
                torch.cuda.synchronize()
                assert np.array_equal(out.cpu().numpy(), want[r0:r1]), (h, w, spread, world, rank)


def test_envelope_cols_writes_only_its_columns(pkg, oracle):
    h, w, spread = 400, 700, 64
    img = oracle.syn_discs(w, h)
    want = oracle.sdf(img, w, h, spread=spread)
    d_img = torch.from_numpy(img).cuda()
    inter = torch.empty((h, pkg.inter_row_bytes(w, spread)), dtype=torch.uint8, device="cuda")
    pkg.rowscan_device(d_img, w, h, spread, inter)
    for (c0, c1) in ((0, 700), (64, 128), (100, 301), (650, 700), (31, 33)):
        out = torch.full((h, w), 0x77, dtype=torch.uint8, device="cuda")
        pkg.envelope_cols_device(inter, w, h, 0, h, c0, c1, spread, False, out)
        torch.cuda.synchronize()
        got = out.cpu().numpy()
        a, b = c0 // 32 * 32, min((c1 + 31) // 32 * 32, w)
        assert np.array_equal(got[:, a:b], want[:, a:b]), (c0, c1)
        assert (got[:, :a] == 0x77).all() and (got[:, b:] == 0x77).all(), (c0, c1)


def test_glyph_atlas_front_end(pkg, oracle):
    """SURVEY.md 8(f) row 4 through the C ABI (chaqsdf_atlas_u8): SYN-GLYPH canvases in, packed atlas out. Every glyph's
    rectangle of the atlas holds the oracle's SDF of that canvas cropped to ink box + spread; everything else is 0."""
    n, side, spread, aw, pad = 96, 256, 16, 2048, 1
    glyphs = oracle.syn_glyphs(n, side)
    glyphs[5] = 0                              # a canvas without ink: no rectangle
    glyphs[6] = 0
    glyphs[6, 0:3, 250:256] = 255              # ink in a corner: the box is clipped to the canvas
    atlas, boxes, xy = pkg.atlas_u8(glyphs, n, side, side, spread=spread, atlas_w=aw, pad=pad)
    assert atlas.shape[1] == aw
    covered = np.zeros(atlas.shape, bool)
    for i in range(n):
        ys, xs = np.nonzero(glyphs[i] > 127)
        x0, y0, x1, y1 = boxes[i].tolist()
        if len(ys) == 0:
            assert (x0, y0, x1, y1) == (0, 0, 0, 0)
            continue
        assert (x0, y0, x1, y1) == (max(xs.min() - spread, 0), max(ys.min() - spread, 0),
                                    min(xs.max() + 1 + spread, side), min(ys.max() + 1 + spread, side)), i
        want = oracle.sdf(glyphs[i], side, side, spread=spread)[y0:y1, x0:x1]
        ax, ay = xy[i].tolist()
        assert np.array_equal(atlas[ay:ay + (y1 - y0), ax:ax + (x1 - x0)], want), i
        assert not covered[ay:ay + (y1 - y0), ax:ax + (x1 - x0)].any(), "rectangles overlap"
        covered[ay:ay + (y1 - y0), ax:ax + (x1 - x0)] = True
    assert (atlas[~covered] == 0).all()
    # the device pieces on their own: boxes with another margin, blit into a caller's atlas
    d = torch.from_numpy(glyphs).cuda()
    b0 = pkg.glyph_boxes_device(d, n, side, side, margin=0).cpu().numpy()
    ys, xs = np.nonzero(glyphs[0] > 127)
    assert b0[0].tolist() == [xs.min(), ys.min(), xs.max() + 1, ys.max() + 1]


@pytest.mark.parametrize("n,spread", [(4096, 64)])
def test_config2_full_size(pkg, oracle, ref, n, spread):
    img = oracle.syn_discs(n)
    got = gpu_u8(pkg, img, n, n, spread=spread)
    assert np.array_equal(got, oracle.sdf(img, n, n, spread=spread))
    assert np.array_equal(got, ref.sdf_u8(img, n, n, spread=spread))  # stock OpenMP build is exact at 4096


def test_config4_full_size(pkg, oracle):
    n, spread = 16384, 256
    img = oracle.syn_discs(n)
    got = gpu_u8(pkg, img, n, n, spread=spread)
    want = oracle.sdf(img, n, n, spread=spread)
    assert np.array_equal(got, want)
    # size-independent property: the transform commutes with transposition
    gt = gpu_u8(pkg, np.ascontiguousarray(img.T), n, n, spread=spread)
    assert np.array_equal(gt, want.T)


def test_large_spread_instances(pkg, oracle):
    """Spreads beyond the banded path (> 256), including BASELINE config 5's spread 1024 on enough rows (>= 2304) that
    its own pass-2 instance runs, on shapes and on noise, symmetric and asymmetric."""
    rng = np.random.default_rng(41)
    w, h = 1536, 2560
    discs = oracle.syn_discs(w, h)
    noise = (rng.random((h, w)) < 1 / 3000).astype(np.uint8) * 255
    for img in (discs, noise):
        for spread, asym in ((1024, False), (1024, True), (300, False), (700, True)):
            got = gpu_u8(pkg, img, w, h, spread=spread, asymmetric=asym)
            assert np.array_equal(got, oracle.sdf(img, w, h, spread=spread, asymmetric=asym)), (spread, asym)
    tall = np.zeros((5000, 96), np.uint8)
    tall[2400:2600, 30:60] = 255
    assert np.array_equal(gpu_u8(pkg, tall, 96, 5000, spread=2000), oracle.sdf(tall, 96, 5000, spread=2000))


def test_unwindowed_paths_on_tall_images(pkg, oracle, ref):
    """The exact-fields entry and the df.h drop-in on images of >= 3072 rows (where the column state no longer fits
    shared memory and lives in the workspace), and a spread larger than the image on that many rows."""
    rng = np.random.default_rng(43)
    n = 3200
    img = random_shapes(rng, n, n)
    res = pkg.sdf_fields_device(torch.from_numpy(img).cuda(), n, n, spread=64)
    torch.cuda.synchronize()
    out, sdf, din, dout = oracle.sdf(img, n, n, spread=64, fields=True)
    assert np.array_equal(res["d2_in"].cpu().numpy(), din) and np.array_equal(res["d2_out"].cpu().numpy(), dout)
    assert np.array_equal(res["out"].cpu().numpy(), out)
    assert np.array_equal(res["sdf"].cpu().numpy().view(np.uint32), sdf.view(np.uint32))
    del res
    assert np.array_equal(gpu_u8(pkg, img, n, n, spread=5000), oracle.sdf(img, n, n, spread=5000))
    n = 4096  # the largest side at which the stock build is exact: 0 ulp against it
    seeds = rng.random((n, n)) < 0.00005
    f = np.where(seeds, np.float32(0), np.float32(np.inf)).astype(np.float32)
    got = pkg.dist_transform_2d(f.copy())
    assert np.array_equal(got.view(np.uint32), ref.dist_transform_2d(f).view(np.uint32))


def test_config5_at_its_shape_band_wise(pkg, oracle):
    """BASELINE config 5 (SYN-DISCS(32768), spread 1024) on one GPU, checked against the oracle band by band: the
    oracle runs on each band plus 1024 context rows (exact for the band's rows: nothing farther away can change a byte).
    Bands at the top edge, in the middle and at the bottom edge."""
    n, spread = 32768, 1024
    img = oracle.syn_discs(n)
    d = torch.from_numpy(img).cuda()
    out = pkg.sdf_u8_device(d, n, n, spread=spread)
    torch.cuda.synchronize()
    got = out.cpu().numpy()
    del d, out
    for (a, b) in ((0, 384), (15000, 15384), (24001, 24300), (n - 384, n)):
        lo, hi = max(0, a - spread - 1), min(n, b + spread + 1)
        want = oracle.sdf(img[lo:hi], n, hi - lo, spread=spread)[a - lo:b - lo]
        assert np.array_equal(got[a:b], want), (a, b)


def test_two_host_threads(pkg, oracle):
    """dist_transform_2d is called from two OpenMP section threads at once by the reference's main (sdfgen.c:277-289);
    chaqsdf_u8 may be too. Two threads, different inputs, many rounds: every result must be the single-threaded one."""
    import threading
    rng = np.random.default_rng(47)
    n = 700
    fields = [np.where(rng.random((n, n)) < p, np.float32(0), np.float32(np.inf)).astype(np.float32) for p in (0.001, 0.01)]
    want_f = [oracle.dist_transform_2d_binary(f) for f in fields]
    imgs = [random_shapes(rng, 1100, 900), oracle.syn_discs(900, 1100)]
    want_b = [oracle.sdf(im, 900, 1100, spread=48) for im in imgs]
    errors = []

    def work(k):
        try:
            for rep in range(12):
                got = pkg.dist_transform_2d(fields[k].copy())
                if not np.array_equal(got.view(np.uint32), want_f[k].view(np.uint32)):
                    errors.append(("dist_transform_2d", k, rep))
                if not np.array_equal(pkg.sdf_u8(imgs[k], 900, 1100, spread=48), want_b[k]):
                    errors.append(("sdf_u8", k, rep))
        except Exception as e:  # noqa: BLE001
            errors.append((k, repr(e)))

    ts = [threading.Thread(target=work, args=(k,)) for k in range(2)]
    for t in ts:
        t.start()
    for t in ts:
        t.join()
    assert not errors, errors[:5]


def test_pass2_organisations_and_run_lengths_agree(pkg, oracle, monkeypatch):
    """Pass 2 with forced run lengths (a junction every 32 rows .. one run per column; junction repair in rounds and
    sequential; four sub-runs per build CTA): all must give the oracle's bytes. The spread 20 / 24 cases with runs of
    3 blocks are the ones in which two junctions of one round would share a hull word if rounds were counted in rows
    instead of words (hull_merge_phases)."""
    rng = np.random.default_rng(77)
    cases = [(oracle.syn_discs(1536), 1536, 1536, 256), (oracle.syn_discs(1024), 1024, 1024, 40),
             (oracle.syn_noise(700), 700, 700, 9), (random_shapes(rng, 900, 333), 333, 900, 100),
             (oracle.syn_noise(1100), 1100, 1100, 20), (random_shapes(rng, 1500, 640), 640, 1500, 24),
             (random_shapes(rng, 2000, 200), 200, 2000, 100)]
    envs = [{}, {"CHAQSDF_HULL_PB": "1", "CHAQSDF_HULL_EB": "1"}, {"CHAQSDF_HULL_PB": "3", "CHAQSDF_HULL_EB": "5"},
            {"CHAQSDF_HULL_PB": "32", "CHAQSDF_HULL_EB": "32"}, {"CHAQSDF_HULL_SEQ": "1", "CHAQSDF_HULL_PB": "2"},
            {"CHAQSDF_HULL_PB": "13", "CHAQSDF_HULL_SUB": "1"}, {"CHAQSDF_HULL_PB": "8", "CHAQSDF_HULL_SUB": "4"}]
    keys = ("CHAQSDF_HULL_PB", "CHAQSDF_HULL_EB", "CHAQSDF_HULL_SEQ", "CHAQSDF_HULL_SUB")
    try:
        for img, w, h, spread in cases:
            want = oracle.sdf(img, w, h, spread=spread)
            for env in envs:
                for k in keys:
                    monkeypatch.delenv(k, raising=False)
                for k, v in env.items():
                    monkeypatch.setenv(k, v)
                pkg.reload_tuning()
                for rep in range(3 if env.get("CHAQSDF_HULL_PB") == "3" else 1):  # a race would not show every time
                    assert np.array_equal(gpu_u8(pkg, img, w, h, spread=spread), want), (w, h, spread, env)
    finally:
        for k in keys:
            monkeypatch.delenv(k, raising=False)
        pkg.reload_tuning()


def test_host_api_pipelined_bands(pkg, oracle, monkeypatch):
    """chaqsdf_u8 on a large image goes through in row bands (upload / kernels / download overlapped); the result
    must not depend on the band height, including bands barely taller than the context they need."""
    n, spread = 2048, 40
    img = oracle.syn_discs(n)
    want = oracle.sdf(img, n, n, spread=spread)
    for rows in ("160", "352", "1000"):
        monkeypatch.setenv("CHAQSDF_PIPE_ROWS", rows)
        pkg.reload_tuning()
        assert np.array_equal(pkg.sdf_u8(img, n, n, spread=spread), want), rows
    # interleaved (Y,A) input, asymmetric, pinned buffers
    ya = np.ascontiguousarray(np.stack([255 - img, img], axis=-1))
    monkeypatch.setenv("CHAQSDF_PIPE_ROWS", "224")
    pkg.reload_tuning()
    hin, hout = torch.from_numpy(ya).pin_memory(), torch.empty((n, n), dtype=torch.uint8).pin_memory()
    pkg.sdf_u8(hin, n, n, stride=2, offset=1, spread=spread, asymmetric=True, out=hout)
    assert np.array_equal(hout.numpy(), oracle.sdf(img, n, n, spread=spread, asymmetric=True))
    monkeypatch.delenv("CHAQSDF_PIPE_ROWS")
    pkg.reload_tuning()
    assert np.array_equal(pkg.sdf_u8(img, n, n, spread=spread), want)
    # repeated calls on alternating images: a stale or early read of the pooled device buffers would show
    flipped = np.ascontiguousarray(img[::-1, ::-1])
    want_f = oracle.sdf(flipped, n, n, spread=spread)
    monkeypatch.setenv("CHAQSDF_PIPE_ROWS", "192")
    pkg.reload_tuning()
    for rep in range(6):
        src, ref_bytes = (img, want) if rep % 2 == 0 else (flipped, want_f)
        assert np.array_equal(pkg.sdf_u8(src, n, n, spread=spread), ref_bytes), rep
    monkeypatch.delenv("CHAQSDF_PIPE_ROWS")
    pkg.reload_tuning()


def test_error_paths(pkg):
    d = torch.zeros((8, 8), dtype=torch.uint8, device="cuda")
    with pytest.raises(pkg.ChaqSdfError):
        pkg.sdf_u8_device(d, 8, 8, spread=0)
    with pytest.raises(pkg.ChaqSdfError):
        pkg.sdf_u8_device(d, 8, 8, spread=4, workspace=torch.zeros(16, dtype=torch.uint8, device="cuda"))


def test_cli_golden_pair_and_flags(pkg, oracle, tmp_path):
    """The host program (reference sdfgen.c CLI contract) on the README example: -s 100 -al on the sample input
    must give the reference's sample output, through PNG files and through stdin/stdout."""
    import subprocess
    from PIL import Image
    cli = os.path.join(ROOT, "chaq-sdfgen_b200", "bin", "chaq_sdfgen")
    if not os.path.exists(cli):
        pytest.skip("bin/chaq_sdfgen not built")
    z = np.load(os.path.join(GOLDEN, "sample_pair.npz"))
    ya = z["ya"].reshape(200, 200, 2)
    src = tmp_path / "in.png"
    Image.fromarray(np.ascontiguousarray(ya[:, :, 0]), "L").save(src)  # 1-channel input: alpha decodes as 255
    dst = tmp_path / "out.png"
    r = subprocess.run([cli, "-i", str(src), "-o", str(dst), "-s", "100", "-al"], capture_output=True, timeout=120)
    assert r.returncode == 0, r.stderr
    assert np.array_equal(np.asarray(Image.open(dst)), z["out"])
    r = subprocess.run([cli, "-i", "-", "-o", "-", "-s", "100", "-la", "-f", "bmp"], input=src.read_bytes(),
                       capture_output=True, timeout=120)
    assert r.returncode == 0 and r.stdout[:2] == b"BM"
    import io
    assert np.array_equal(np.asarray(Image.open(io.BytesIO(r.stdout)).convert("L")), z["out"])
    # default test is the alpha channel: an image without alpha is all inside -> all 255
    r = subprocess.run([cli, "-i", str(src), "-o", str(dst), "-s", "8"], capture_output=True, timeout=120)
    assert r.returncode == 0 and (np.asarray(Image.open(dst)) == 255).all()
    # -n on luminance, RGBA input with a real alpha channel
    rng = np.random.default_rng(3)
    rgba = rng.integers(0, 256, (90, 130, 4), dtype=np.uint8)
    rgba[20:60, 30:100] = (250, 250, 250, 255)
    Image.fromarray(rgba, "RGBA").save(src)
    c32 = rgba.astype(np.int32)
    lum = ((77 * c32[..., 0] + 150 * c32[..., 1] + 29 * c32[..., 2]) >> 8).astype(np.uint8)  # stb_image.h:1726-1729
    for flags, img, above in ((["-ln"], lum, False), ([], rgba[..., 3], True), (["-a"], rgba[..., 3], True)):
        r = subprocess.run([cli, "-i", str(src), "-o", str(dst), "-s", "12", *flags], capture_output=True, timeout=120)
        assert r.returncode == 0, r.stderr
        want = oracle.sdf(np.ascontiguousarray(img), 130, 90, spread=12, test_above=above, asymmetric="-a" in flags)
        assert np.array_equal(np.asarray(Image.open(dst)), want), flags


def test_cli_pgm_passthrough(pkg, oracle, tmp_path):
    """Binary PGM in and out (for images beyond stb_image's reach): header with a comment, luminance test, stdin."""
    import subprocess
    cli = os.path.join(ROOT, "chaq-sdfgen_b200", "bin", "chaq_sdfgen")
    if not os.path.exists(cli):
        pytest.skip("bin/chaq_sdfgen not built")
    w, h, spread = 300, 170, 21
    img = oracle.syn_discs(w, h)
    src, dst = tmp_path / "in.pgm", tmp_path / "out.pgm"
    src.write_bytes(b"P5\n# made by a test\n%d %d\n255\n" % (w, h) + img.tobytes())
    want = oracle.sdf(img, w, h, spread=spread)

    def body(raw):
        assert raw.startswith(b"P5\n%d %d\n255\n" % (w, h))
        return np.frombuffer(raw[len(b"P5\n%d %d\n255\n" % (w, h)):], np.uint8).reshape(h, w)

    r = subprocess.run([cli, "-i", str(src), "-o", str(dst), "-s", str(spread), "-l"], capture_output=True, timeout=120)
    assert r.returncode == 0, r.stderr
    assert np.array_equal(body(dst.read_bytes()), want)
    r = subprocess.run([cli, "-i", "-", "-o", "-", "-f", "pgm", "-s", str(spread), "-l"], input=src.read_bytes(),
                       capture_output=True, timeout=120)
    assert r.returncode == 0 and np.array_equal(body(r.stdout), want)
    # no alpha channel in a PGM: the default (alpha) test sees an opaque image -> all inside -> all 255
    r = subprocess.run([cli, "-i", str(src), "-o", str(dst), "-s", str(spread)], capture_output=True, timeout=120)
    assert r.returncode == 0 and (body(dst.read_bytes()) == 255).all()


def test_narrow_rows_several_per_warp(pkg, oracle):
    """Rows of <= 512 pixels go several to a warp in pass 1 (rowscan8_narrow_kernel): widths on both sides of every
    lanes-per-row step, aligned (128-bit loads) and not, strided, row counts that leave a warp partly empty, batches."""
    rng = np.random.default_rng(77)
    for (h, w) in [(5, 16), (9, 48), (3, 64), (67, 96), (13, 128), (70, 144), (41, 256), (19, 272), (33, 496), (21, 512),
                   (11, 500), (6, 40)]:
        for img in (random_shapes(rng, h, w), (rng.random((h, w)) < 0.5).astype(np.uint8) * 255,
                    rng.integers(0, 256, (h, w), dtype=np.uint8)):
            for spread in (2, 20, 256):
                got = gpu_u8(pkg, img, w, h, spread=spread)
                assert np.array_equal(got, oracle.sdf(img, w, h, spread=spread)), (h, w, spread)
    h, w = 37, 208
    ya = rng.integers(0, 256, (h, w, 2), dtype=np.uint8)
    ya[5:30, 20:150, :] = 220
    for off in (0, 1):
        for above in (True, False):
            got = gpu_u8(pkg, ya, w, h, stride=2, offset=off, test_above=above, spread=7)
            assert np.array_equal(got, oracle.sdf(ya, w, h, stride=2, offset=off, test_above=above, spread=7))
    n, s = 5, 48   # 240 rows of 2 words: the last warp of the grid is partly empty
    imgs = np.stack([random_shapes(rng, s, s) for _ in range(n)])
    out = pkg.sdf_u8_batched_device(torch.from_numpy(imgs).cuda(), n, s, s, spread=16)
    torch.cuda.synchronize()
    got = out.cpu().numpy()
    for k in range(n):
        assert np.array_equal(got[k], oracle.sdf(imgs[k], s, s, spread=16)), k


def test_multi_device_entry(pkg, oracle):
    """chaqsdf_u8_multi: row bands over a list of devices of this process (host buffers in and out), byte-identical to
    the oracle. A device may be listed more than once -- every position is a band with its own buffers -- so one GPU
    checks the band logic too: halos from one neighbour, from several (spread larger than a band), the general
    intermediate (spread > 256), strided input, more devices than rows."""
    rng = np.random.default_rng(53)
    ngpu = torch.cuda.device_count()
    lists = [[0], [0, 0], [0, 0, 0, 0, 0]] + ([list(range(ngpu)), list(range(ngpu)) * 2] if ngpu > 1 else [])
    w, h = 640, 900
    img = oracle.syn_discs(w, h)
    for devs in lists:
        for spread, asym in ((40, False), (256, True), (300, False), (1000, False)):
            got = pkg.sdf_u8_multi(img, w, h, devs, spread=spread, asymmetric=asym)
            assert np.array_equal(got, oracle.sdf(img, w, h, spread=spread, asymmetric=asym)), (devs, spread)
    ya = rng.integers(0, 256, (333, 500, 2), dtype=np.uint8)
    ya[100:200, 50:400] = (3, 250)
    got = pkg.sdf_u8_multi(ya, 500, 333, [0, 0, 0], stride=2, offset=1, test_above=False, spread=20)
    assert np.array_equal(got, oracle.sdf(ya, 500, 333, stride=2, offset=1, test_above=False, spread=20))
    tiny = random_shapes(rng, 3, 70)
    assert np.array_equal(pkg.sdf_u8_multi(tiny, 70, 3, [0] * 8, spread=5), oracle.sdf(tiny, 70, 3, spread=5))
    big = oracle.syn_discs(4096)  # large enough for every band to be a real launch
    want = oracle.sdf(big, 4096, 4096, spread=64)
    for devs in lists[1:]:
        assert np.array_equal(pkg.sdf_u8_multi(big, 4096, 4096, devs, spread=64), want), devs
    with pytest.raises(pkg.ChaqSdfError):
        pkg.sdf_u8_multi(img, w, h, [ngpu], spread=8)  # no such device
    assert torch.cuda.current_device() == 0


def test_cli_gpus_flag_and_16bit_pgm(pkg, oracle, tmp_path):
    """-g n (row bands over the first n GPUs, chaqsdf_u8_multi) gives the same file as the single-GPU run; a 16-bit PGM
    is not for the private 8-bit reader and goes to stb_image like every other format (as the reference would)."""
    import subprocess
    cli = os.path.join(ROOT, "chaq-sdfgen_b200", "bin", "chaq_sdfgen")
    if not os.path.exists(cli):
        pytest.skip("bin/chaq_sdfgen not built")
    n = 700
    img = oracle.syn_discs(n)
    src, a, b = tmp_path / "in.pgm", tmp_path / "a.pgm", tmp_path / "b.pgm"
    src.write_bytes(b"P5\n%d %d\n255\n" % (n, n) + img.tobytes())
    g = min(torch.cuda.device_count(), 4)
    r = subprocess.run([cli, "-i", str(src), "-o", str(a), "-s", "30", "-l"], capture_output=True, timeout=120)
    assert r.returncode == 0, r.stderr
    r = subprocess.run([cli, "-i", str(src), "-o", str(b), "-s", "30", "-l", "-g", str(g), "-t"], capture_output=True, timeout=120)
    assert r.returncode == 0 and b"GPU(s)" in r.stderr, r.stderr
    assert a.read_bytes() == b.read_bytes()
    assert np.array_equal(np.frombuffer(a.read_bytes()[-n * n:], np.uint8).reshape(n, n), oracle.sdf(img, n, n, spread=30))
    r = subprocess.run([cli, "-i", str(src), "-o", str(b), "-g", str(torch.cuda.device_count() + 1)], capture_output=True, timeout=120)
    assert r.returncode == 255 and b"device index out of range" in r.stderr
    r = subprocess.run([cli, "-i", str(src), "-o", str(b), "-g", "0"], capture_output=True, timeout=120)
    assert r.returncode == 255 and b"Invalid value given for gpus" in r.stderr
    # a PGM with a lower maxval is stb_image's (raw bytes, as in the reference); a 16-bit one is refused with a message
    small = np.ascontiguousarray(img[:90, :120]) // 2  # values 0 / 127: with -l (> 127) nothing is inside
    small[10:40, 20:70] = 200
    low = tmp_path / "low.pgm"
    low.write_bytes(b"P5\n120 90\n200\n" + small.tobytes())
    r = subprocess.run([cli, "-i", str(low), "-o", str(a), "-s", "9", "-l", "-f", "pgm"], capture_output=True, timeout=120)
    assert r.returncode == 0, r.stderr
    got = np.frombuffer(a.read_bytes()[-90 * 120:], np.uint8).reshape(90, 120)
    assert np.array_equal(got, oracle.sdf(small, 120, 90, spread=9))
    r = subprocess.run([cli, "-i", "-", "-o", "-", "-s", "9", "-l", "-f", "pgm"], input=low.read_bytes(), capture_output=True, timeout=120)
    assert r.returncode == 0 and np.array_equal(np.frombuffer(r.stdout[-90 * 120:], np.uint8).reshape(90, 120), got)
    wide = (small.astype(np.uint16) * 257).astype(">u2")
    src16 = tmp_path / "in16.pgm"
    src16.write_bytes(b"P5\n120 90\n65535\n" + wide.tobytes())
    r = subprocess.run([cli, "-i", str(src16), "-o", str(a), "-s", "9", "-l"], capture_output=True, timeout=120)
    assert r.returncode == 255 and b"16-bit PGM" in r.stderr


def test_reference_main_runs_on_the_dropin(pkg, tmp_path):
    """The reference's own main (openmp/sdfgen.c, unmodified) linked against libchaqsdf_b200.so instead of df.c
    (oracle/Makefile builds oracle/_ref/chaq_sdfgen_dropin where the reference tree exists):
    the README example through it must give the reference's sample output. Its two OpenMP section threads call
    dist_transform_2d at the same time (sdfgen.c:277-289)."""
    import subprocess
    from PIL import Image
    exe = os.path.join(ROOT, "oracle", "_ref", "chaq_sdfgen_dropin")
    if not os.path.exists(exe):
        pytest.skip("oracle/_ref/chaq_sdfgen_dropin not built (needs the reference tree)")
    z = np.load(os.path.join(GOLDEN, "sample_pair.npz"))
    ya = z["ya"].reshape(200, 200, 2)
    src, dst = tmp_path / "in.png", tmp_path / "out.png"
    Image.fromarray(np.ascontiguousarray(ya[:, :, 0]), "L").save(src)
    for _ in range(3):
        r = subprocess.run([exe, "-i", str(src), "-o", str(dst), "-s", "100", "-al"], capture_output=True, timeout=300)
        assert r.returncode == 0, r.stderr
        assert np.array_equal(np.asarray(Image.open(dst)), z["out"])
    big = tmp_path / "big.png"  # larger, non-trivial alpha test (default flags), both section threads busy for a while
    rng = np.random.default_rng(5)
    rgba = np.zeros((1500, 1500, 4), np.uint8)
    rgba[..., 3] = random_shapes(rng, 1500, 1500)
    Image.fromarray(rgba, "RGBA").save(big)
    r = subprocess.run([exe, "-i", str(big), "-o", str(dst), "-s", "32"], capture_output=True, timeout=300)
    assert r.returncode == 0, r.stderr
    from oracle import Oracle
    assert np.array_equal(np.asarray(Image.open(dst)), Oracle().sdf(np.ascontiguousarray(rgba[..., 3]), 1500, 1500, spread=32))


def test_exact_fields_by_column_parts(pkg, oracle):
    """The pieces of the multi-GPU exact path (row bands -> all-to-all -> column parts, SURVEY.md 8(e) variant E1) on
    one device: pass 1 of two row bands scattered into three column-part buffers (the store is the all-to-all), pass 2
    per part as an image of the part's width; the parts side by side must be the oracle's exact fields."""
    rng = np.random.default_rng(61)
    for (h, w, nparts, bands) in ((300, 1000, 3, 2), (257, 96, 4, 3), (64, 2050, 8, 1), (1100, 700, 2, 5)):
        img = random_shapes(rng, h, w)
        d = torch.from_numpy(img).cuda()
        parts = []
        for q in range(nparts):
            c0, c1 = pkg.col_part(w, nparts, q)
            parts.append(torch.zeros((h, max(c1 - c0, 1)), dtype=torch.int16, device="cuda"))
        for b in range(bands):
            r0, r1 = pkg.shard_rows(h, bands, b)
            pkg.rowscan_scatter_device(d[r0:r1], w, r1 - r0, r0, [p.data_ptr() for p in parts])
        out, sdf, din, dout = oracle.sdf(img, w, h, spread=24, fields=True)
        for q in range(nparts):
            c0, c1 = pkg.col_part(w, nparts, q)
            if c1 == c0:
                continue
            res = pkg.fields_from_inter_device(parts[q][:, :c1 - c0].contiguous() if parts[q].shape[1] != c1 - c0 else parts[q],
                                               c1 - c0, h, spread=24, want=("d2_in", "d2_out", "out", "sdf"))
            torch.cuda.synchronize()
            assert np.array_equal(res["d2_in"].cpu().numpy(), din[:, c0:c1]), (h, w, q)
            assert np.array_equal(res["d2_out"].cpu().numpy(), dout[:, c0:c1]), (h, w, q)
            assert np.array_equal(res["out"].cpu().numpy(), out[:, c0:c1]), (h, w, q)
            assert np.array_equal(res["sdf"].cpu().numpy().view(np.uint32), sdf[:, c0:c1].view(np.uint32)), (h, w, q)
    # world 1 through the sharded entry (no process group needed)
    img = random_shapes(rng, 500, 640)
    res = pkg.sdf_fields_a2a(torch.from_numpy(img).cuda(), 640, 500, spread=24)
    out, sdf, din, dout = oracle.sdf(img, 640, 500, spread=24, fields=True)
    assert np.array_equal(res["d2_in"].cpu().numpy(), din) and np.array_equal(res["out"].cpu().numpy(), out)
