"""ctypes mirror of include/chaqsdf.h. Names follow the reference: dist_transform_2d (openmp/df.h:6) and the
sdfgen.c pipeline (openmp/sdfgen.c:261-300) as sdf_u8*. Device functions take torch CUDA tensors."""
import ctypes as C
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("CHAQSDF_LIB") or os.path.join(HERE, "libchaqsdf_b200.so")  # override: A/B builds only
_lib = None


class ChaqSdfError(RuntimeError):
    pass


def build_library(force=False):
    """nvcc -gencode arch=compute_100a,code=sm_100a (csrc/Makefile). Cross-compiles without a GPU."""
    if force and os.path.exists(LIB_PATH):
        os.remove(LIB_PATH)
    r = subprocess.run(["make", "-C", os.path.join(HERE, "csrc"), "-s"], capture_output=True, text=True)
    if r.returncode != 0:
        raise ChaqSdfError("building libchaqsdf_b200.so failed:\n" + r.stdout + r.stderr)
    return LIB_PATH


SYMBOLS = ["chaqsdf_version", "chaqsdf_last_error_string", "chaqsdf_launch_count", "chaqsdf_reload_tuning",
           "chaqsdf_profile_enable",
           "chaqsdf_profile_mean_ms", "dist_transform_2d",
           "chaqsdf_dt2d_last_status", "chaqsdf_u8", "chaqsdf_u8_multi", "chaqsdf_pipeline_plan", "chaqsdf_workspace_bytes", "chaqsdf_u8_device",
           "chaqsdf_u8_device_static", "chaqsdf_envelope_cols_device", "chaqsdf_transpose_u8_device",
           "chaqsdf_atlas_pack", "chaqsdf_glyph_boxes_device", "chaqsdf_atlas_blit_device", "chaqsdf_atlas_u8",
           "chaqsdf_batched_workspace_bytes", "chaqsdf_u8_batched_device", "chaqsdf_rowscan_device",
           "chaqsdf_inter_row_bytes",
           "chaqsdf_envelope_workspace_bytes", "chaqsdf_envelope_device", "chaqsdf_fields_workspace_bytes",
           "chaqsdf_fields_device", "chaqsdf_col_part", "chaqsdf_rowscan_scatter_device",
           "chaqsdf_fields_from_inter_workspace_bytes", "chaqsdf_fields_from_inter_device"]


def lib():
    """Load libchaqsdf_b200.so. No fallback: a missing library is an error."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise ChaqSdfError(f"{LIB_PATH} not built: run __graft_entry__.build() (there is no CPU fallback)")
    L = C.CDLL(LIB_PATH)
    sz, vp, i = C.c_size_t, C.c_void_p, C.c_int
    L.chaqsdf_last_error_string.restype = C.c_char_p
    L.chaqsdf_launch_count.restype = C.c_ulonglong
    L.chaqsdf_profile_enable.argtypes = [i]
    L.chaqsdf_profile_enable.restype = None
    L.chaqsdf_reload_tuning.restype = None
    L.chaqsdf_profile_mean_ms.argtypes = [C.POINTER(C.c_float)]
    L.dist_transform_2d.argtypes = [vp, sz, sz]
    L.dist_transform_2d.restype = None
    L.chaqsdf_u8.argtypes = [vp, sz, sz, sz, sz, i, sz, i, vp]
    L.chaqsdf_u8_multi.argtypes = [vp, sz, sz, sz, sz, i, sz, i, C.POINTER(C.c_int), i, vp]
    L.chaqsdf_pipeline_plan.argtypes = [sz, sz, sz, sz, C.POINTER(sz), sz]
    L.chaqsdf_pipeline_plan.restype = sz
    L.chaqsdf_workspace_bytes.argtypes = [sz, sz, sz]
    L.chaqsdf_workspace_bytes.restype = sz
    L.chaqsdf_batched_workspace_bytes.argtypes = [sz, sz, sz, sz]
    L.chaqsdf_batched_workspace_bytes.restype = sz
    L.chaqsdf_envelope_workspace_bytes.argtypes = [sz, sz, sz]
    L.chaqsdf_envelope_workspace_bytes.restype = sz
    L.chaqsdf_inter_row_bytes.argtypes = [sz, sz]
    L.chaqsdf_inter_row_bytes.restype = sz
    L.chaqsdf_fields_workspace_bytes.argtypes = [sz, sz]
    L.chaqsdf_fields_workspace_bytes.restype = sz
    L.chaqsdf_u8_device.argtypes = [vp, sz, sz, sz, sz, i, sz, i, vp, vp, sz, vp]
    L.chaqsdf_u8_device_static.argtypes = [vp, sz, sz, sz, sz, i, sz, i, vp, vp]
    L.chaqsdf_u8_batched_device.argtypes = [vp, sz, sz, sz, sz, sz, i, sz, i, vp, vp, sz, vp]
    L.chaqsdf_rowscan_device.argtypes = [vp, sz, sz, sz, sz, i, sz, vp, vp]
    L.chaqsdf_envelope_device.argtypes = [vp, sz, sz, sz, sz, sz, i, vp, vp, sz, vp]
    L.chaqsdf_envelope_cols_device.argtypes = [vp, sz, sz, sz, sz, sz, sz, sz, i, vp, vp, sz, vp]
    L.chaqsdf_transpose_u8_device.argtypes = [vp, sz, sz, sz, vp, sz, vp]
    L.chaqsdf_atlas_pack.argtypes = [sz, vp, sz, sz, vp, C.POINTER(sz)]
    L.chaqsdf_glyph_boxes_device.argtypes = [vp, sz, sz, sz, i, sz, vp, vp]
    L.chaqsdf_atlas_blit_device.argtypes = [vp, sz, sz, sz, vp, vp, vp, sz, vp]
    L.chaqsdf_atlas_u8.argtypes = [vp, sz, sz, sz, i, sz, i, sz, sz, vp, sz, C.POINTER(sz), vp, vp]
    L.chaqsdf_fields_device.argtypes = [vp, sz, sz, sz, sz, i, sz, i, vp, vp, vp, vp, vp, vp, vp, sz, vp]
    L.chaqsdf_col_part.argtypes = [sz, sz, sz, C.POINTER(sz), C.POINTER(sz)]
    L.chaqsdf_col_part.restype = sz
    L.chaqsdf_rowscan_scatter_device.argtypes = [vp, sz, sz, sz, sz, i, sz, sz, C.POINTER(vp), vp]
    L.chaqsdf_fields_from_inter_workspace_bytes.argtypes = [sz, sz]
    L.chaqsdf_fields_from_inter_workspace_bytes.restype = sz
    L.chaqsdf_fields_from_inter_device.argtypes = [vp, sz, sz, sz, i, vp, vp, vp, vp, vp, vp, vp, sz, vp]
    _lib = L
    return L


def _check(rc):
    if rc != 0:
        raise ChaqSdfError(f"chaqsdf error {rc}: {lib().chaqsdf_last_error_string().decode()}")


def launch_count():
    return int(lib().chaqsdf_launch_count())


def reload_tuning():
    """Re-read the CHAQSDF_* tuning switches from the environment (they are otherwise read once, at load)."""
    lib().chaqsdf_reload_tuning()


def profile_enable(on=True):
    """Bracket the kernels of the banded pass 2 with CUDA events (bench.py's roofline)."""
    lib().chaqsdf_profile_enable(int(bool(on)))


def profile_mean_ms():
    """{kernel: mean ms} over the profiled calls of the banded pass 2: hull build, junction merge, hull emit."""
    ms = (C.c_float * 3)()
    n = lib().chaqsdf_profile_mean_ms(ms)
    if n <= 0:
        raise ChaqSdfError(lib().chaqsdf_last_error_string().decode())
    return {"hull_build_kernel": float(ms[0]), "hull_merge_par_kernel": float(ms[1]), "hull_paint_kernel": float(ms[2]),
            "calls": int(n)}


def _stream_ptr(stream):
    import torch
    s = stream if stream is not None else torch.cuda.current_stream()
    return C.c_void_p(s.cuda_stream)


def _ptr(t):
    return C.c_void_p(t.data_ptr()) if t is not None else None


# ---------------------------------------------------------------- host-buffer entry points
def dist_transform_2d(img):
    """reference openmp/df.h:6: in place on a C-contiguous float32 numpy array [h, w] of {0, +inf}."""
    assert img.dtype == np.float32 and img.flags.c_contiguous and img.ndim == 2
    h, w = img.shape
    L = lib()
    L.dist_transform_2d(img.ctypes.data_as(C.c_void_p), w, h)
    _check(L.chaqsdf_dt2d_last_status())
    return img


def sdf_u8(img, w, h, stride=1, offset=0, test_above=True, spread=64, asymmetric=False, out=None):
    """reference openmp/sdfgen.c:261-300 on host buffers (numpy, or pinned torch CPU tensors)."""
    if out is None:
        out = np.empty((h, w), np.uint8)
    ip = img.ctypes.data_as(C.c_void_p) if isinstance(img, np.ndarray) else _ptr(img)
    op = out.ctypes.data_as(C.c_void_p) if isinstance(out, np.ndarray) else _ptr(out)
    if isinstance(img, np.ndarray):
        assert img.dtype == np.uint8 and img.flags.c_contiguous and img.size >= w * h * stride
    _check(lib().chaqsdf_u8(ip, w, h, stride, offset, int(test_above), spread, int(asymmetric), op))
    return out


# ---------------------------------------------------------------- device entry points (torch tensors)
def sdf_u8_multi(img, w, h, devices, stride=1, offset=0, test_above=True, spread=64, asymmetric=False, out=None):
    """Host buffers in and out, row bands over the listed devices of this process (chaqsdf_u8_multi)."""
    import numpy as np
    src = np.ascontiguousarray(img, np.uint8)
    if out is None:
        out = np.empty((h, w), np.uint8)
    devs = (C.c_int * len(devices))(*devices)
    _check(lib().chaqsdf_u8_multi(src.ctypes.data_as(C.c_void_p), w, h, stride, offset, int(test_above), spread,
                                  int(asymmetric), devs, len(devices), out.ctypes.data_as(C.c_void_p)))
    return out


def pipeline_plan(w, h, stride=1, spread=64):
    """Band heights (rows, top to bottom) chaqsdf_u8 uses for a host image of this shape; [] = one shot."""
    buf = (C.c_size_t * 4096)()
    n = int(lib().chaqsdf_pipeline_plan(w, h, stride, spread, buf, 4096))
    return [int(buf[k]) for k in range(min(n, 4096))]


def workspace_bytes(w, h, spread, n=1):
    return int(lib().chaqsdf_batched_workspace_bytes(n, w, h, spread))


def _workspace(nbytes, device):
    import torch
    return torch.empty(max(nbytes, 256), dtype=torch.uint8, device=device)


def sdf_u8_device(d_img, w, h, stride=1, offset=0, test_above=True, spread=64, asymmetric=False, out=None,
                  workspace=None, stream=None):
    import torch
    if out is None:
        out = torch.empty((h, w), dtype=torch.uint8, device=d_img.device)
    if workspace is None:
        workspace = _workspace(workspace_bytes(w, h, spread), d_img.device)
    _check(lib().chaqsdf_u8_device(_ptr(d_img), w, h, stride, offset, int(test_above), spread, int(asymmetric),
                                   _ptr(out), _ptr(workspace), workspace.numel(), _stream_ptr(stream)))
    return out


STATIC_SCRATCH_BYTES = 8388608  # CHAQSDF_STATIC_SCRATCH_BYTES of include/chaqsdf.h


def sdf_u8_device_static(d_img, w, h, out, stride=1, offset=0, test_above=True, spread=64, asymmetric=False, stream=None):
    """chaqsdf_u8_device_static: workspace and tables in the library's static device array; allocates nothing
    (the caller passes `out`). Calls that overlap on the GPU share that array and race."""
    _check(lib().chaqsdf_u8_device_static(_ptr(d_img), w, h, stride, offset, int(test_above), spread, int(asymmetric),
                                          _ptr(out), _stream_ptr(stream)))
    return out


def sdf_u8_batched_device(d_imgs, n, w, h, stride=1, offset=0, test_above=True, spread=16, asymmetric=False, out=None,
                          workspace=None, stream=None):
    import torch
    if out is None:
        out = torch.empty((n, h, w), dtype=torch.uint8, device=d_imgs.device)
    if workspace is None:
        workspace = _workspace(workspace_bytes(w, h, spread, n), d_imgs.device)
    _check(lib().chaqsdf_u8_batched_device(_ptr(d_imgs), n, w, h, stride, offset, int(test_above), spread,
                                           int(asymmetric), _ptr(out), _ptr(workspace), workspace.numel(),
                                           _stream_ptr(stream)))
    return out


def inter_row_bytes(w, spread):
    """Bytes per image row of the opaque intermediate between the two passes."""
    return int(lib().chaqsdf_inter_row_bytes(w, spread))


def rowscan_device(d_img, w, rows, spread, d_inter, stride=1, offset=0, test_above=True, stream=None):
    _check(lib().chaqsdf_rowscan_device(_ptr(d_img), w, rows, stride, offset, int(test_above), spread, _ptr(d_inter),
                                        _stream_ptr(stream)))
    return d_inter


def envelope_device(d_inter, w, buf_rows, out_begin, out_end, spread, asymmetric, d_out, workspace=None, stream=None):
    nb = int(lib().chaqsdf_envelope_workspace_bytes(w, buf_rows, spread))
    if workspace is None and nb:
        workspace = _workspace(nb, d_inter.device)
    _check(lib().chaqsdf_envelope_device(_ptr(d_inter), w, buf_rows, out_begin, out_end, spread, int(asymmetric),
                                         _ptr(d_out), _ptr(workspace), workspace.numel() if workspace is not None else 0,
                                         _stream_ptr(stream)))
    return d_out


def sdf_fields_device(d_img, w, h, stride=1, offset=0, test_above=True, spread=64, asymmetric=False, want=("d2_in", "d2_out", "f_in", "f_out", "sdf", "out"),
                      stream=None):
    """Exact unwindowed fields (see chaqsdf_fields_device). Returns a dict of torch tensors."""
    import torch
    dev = d_img.device
    dt = {"d2_in": torch.int32, "d2_out": torch.int32, "f_in": torch.float32, "f_out": torch.float32,
          "sdf": torch.float32, "out": torch.uint8}
    res = {k: torch.empty((h, w), dtype=dt[k], device=dev) for k in want}
    ws = _workspace(int(lib().chaqsdf_fields_workspace_bytes(w, h)), dev)
    g = lambda k: _ptr(res.get(k))  # noqa: E731
    _check(lib().chaqsdf_fields_device(_ptr(d_img), w, h, stride, offset, int(test_above), spread, int(asymmetric),
                                       g("d2_in"), g("d2_out"), g("f_in"), g("f_out"), g("sdf"), g("out"), _ptr(ws),
                                       ws.numel(), _stream_ptr(stream)))
    return res


# ---------------------------------------------------------------- row-band sharding (multi-GPU)
def shard_rows(h, world, rank):
    """Rows [r0, r1) of an h-row image owned by `rank` (contiguous bands, remainder to the first ranks)."""
    base, rem = divmod(h, world)
    r0 = rank * base + min(rank, rem)
    return r0, r0 + base + (1 if rank < rem else 0)


def halo_plan(h, world, rank, spread):
    """Which rows of the 16-bit intermediate this rank receives from / sends to which peer so that its pass 2
    sees `spread` rows of context (or the image edge) around its band. Rows are image rows.
    Returns (r0, r1, halo_top, halo_bot, recvs, sends) with recvs/sends = [(peer, a, b)], rows [a, b)."""
    r0, r1 = shard_rows(h, world, rank)
    halo_top, halo_bot = min(spread, r0), min(spread, h - r1)
    recvs, sends = [], []
    for other in range(world):
        if other == rank:
            continue
        o0, o1 = shard_rows(h, world, other)
        for (a, b) in ((max(o0, r0 - halo_top), min(o1, r0)), (max(o0, r1), min(o1, r1 + halo_bot))):
            if a < b:
                recvs.append((other, a, b))
        oh_top, oh_bot = min(spread, o0), min(spread, h - o1)
        for (a, b) in ((max(r0, o0 - oh_top), min(r1, o0)), (max(r0, o1), min(r1, o1 + oh_bot))):
            if a < b:
                sends.append((other, a, b))
    return r0, r1, halo_top, halo_bot, recvs, sends


def exchange_halos(inter, h, world, rank, spread, group=None):
    """inter: [halo_top + rows + halo_bot, w] tensor whose middle rows are filled; fills the halo rows from
    the neighbouring ranks with batched point-to-point ops (NCCL on GPU tensors, gloo on CPU tensors)."""
    import torch.distributed as dist
    r0, r1, halo_top, halo_bot, recvs, sends = halo_plan(h, world, rank, spread)
    base = r0 - halo_top
    ops = [dist.P2POp(dist.irecv, inter[a - base:b - base], peer, group) for (peer, a, b) in recvs]
    ops += [dist.P2POp(dist.isend, inter[a - base:b - base], peer, group) for (peer, a, b) in sends]
    if ops:
        for req in dist.batch_isend_irecv(ops):
            req.wait()
    return inter


def _peer_setup(state, w, h, spread, rank, world, group, dev):
    """Symmetric (peer-mapped) intermediates for the halo pull: every rank allocates the same [2][rows][rb] buffer
    (two slots, alternating by call), maps its neighbours' copies over NVLink, and all ranks agree on whether that
    worked. Returns True when the peer path is usable everywhere."""
    import torch
    import torch.distributed as dist
    ok = 1
    try:
        import torch.distributed._symmetric_memory as symm
        plans = [halo_plan(h, world, r, spread) for r in range(world)]
        max_rows = max(p[2] + (p[1] - p[0]) + p[3] for p in plans)
        rb = inter_row_bytes(w, spread)
        sym = symm.empty((2, max_rows, rb), dtype=torch.uint8, device=dev)
        hdl = symm.rendezvous(sym, group if group is not None else dist.group.WORLD)
        bufs = {peer: hdl.get_buffer(peer, (2, max_rows, rb), torch.uint8) for (peer, _, _) in plans[rank][4]}
        state.update(sym=sym, hdl=hdl, plans=plans, peer_bufs=bufs, step=0)
    except Exception as e:  # no symmetric memory on this box / build: every rank falls back together
        ok = 0
        state["peer_error"] = repr(e)
    flag = torch.tensor([ok], dtype=torch.int32, device=dev)
    dist.all_reduce(flag, op=dist.ReduceOp.MIN, group=group)
    return bool(flag.item())


def halo_pull_list(plans, rank):
    """The peer-memory exchange of one rank, from the halo plans of all ranks: (signal_to, wait_for, copies) with
    copies = [(peer, dst_lo, dst_hi, src_lo, src_hi)] in buffer rows (dst in this rank's buffer, src in the peer's).
    Only rows a peer OWNS are ever read from it, so a pull never depends on another pull."""
    r0, r1, halo_top, halo_bot, recvs, sends = plans[rank]
    base = r0 - halo_top
    copies = []
    for (peer, a, b) in recvs:
        pbase = plans[peer][0] - plans[peer][2]  # image row held by row 0 of the peer's buffer
        copies.append((peer, a - base, b - base, a - pbase, b - pbase))
    return sorted({p for (p, _, _) in sends}), sorted({p for (p, _, _) in recvs}), copies


def pull_halos(state, slot, rank, timeout_ms=20000):
    """Peer-memory halo exchange: tell the neighbours this rank's rows of slot `slot` are written, wait for theirs,
    then copy the halo rows straight out of their buffers over NVLink. Stream-ordered; no host synchronisation.
    The signals pair up one to one (I signal exactly the ranks that wait for me), and the two slots make the next
    call's pass 1 safe: it overwrites a slot only after the neighbours have signalled the call in between."""
    hdl = state["hdl"]
    signal_to, wait_for, copies = halo_pull_list(state["plans"], rank)
    inter = state["sym"][slot]
    for peer in signal_to:
        hdl.put_signal(peer, slot, timeout_ms)
    for peer in wait_for:
        hdl.wait_signal(peer, slot, timeout_ms)
    for (peer, d0, d1, s0, s1) in copies:
        inter[d0:d1].copy_(state["peer_bufs"][peer][slot][s0:s1], non_blocking=True)


def sdf_u8_sharded(d_band, w, h, spread, asymmetric=False, rank=0, world=1, group=None, state=None, stride=1, offset=0,
                   test_above=True, events=None, halo=None):
    """One image of h rows sharded by row bands over `world` ranks (one process per GPU). d_band holds this
    rank's rows. Pass 1 is local; pass 2 needs `spread` rows of the intermediate from each neighbour band: pulled
    out of the neighbours' peer-mapped buffers over NVLink after a pairwise device-side signal (halo "peer", the
    default on GPUs), or exchanged point to point through NCCL / gloo (halo "p2p"; also the fallback when symmetric
    memory cannot be set up; env CHAQSDF_HALO overrides). Then pass 2 is local. Returns this rank's output rows.
    `state` (a dict) caches buffers between calls; state["halo"] says which exchange is in use."""
    import torch
    r0, r1, halo_top, halo_bot, _, _ = halo_plan(h, world, rank, spread)
    rows = r1 - r0
    buf_rows = halo_top + rows + halo_bot
    dev = d_band.device
    if state is None:
        state = {}
    key = (w, h, spread, rank, world)
    if state.get("key") != key:
        state.clear()
        state["key"] = key
        mode = halo or os.environ.get("CHAQSDF_HALO") or ("peer" if dev.type == "cuda" and world > 1 else "p2p")
        if mode == "peer" and not (world > 1 and _peer_setup(state, w, h, spread, rank, world, group, dev)):
            mode = "p2p"
        state["halo"] = mode
        if mode != "peer":
            state["inter"] = torch.empty((buf_rows, inter_row_bytes(w, spread)), dtype=torch.uint8, device=dev)
        state["out"] = torch.empty((rows, w), dtype=torch.uint8, device=dev)
        nb = int(lib().chaqsdf_envelope_workspace_bytes(w, buf_rows, spread))
        state["ws"] = _workspace(nb, dev) if nb else None
    out = state["out"]
    if state["halo"] == "peer":
        slot = state["step"] & 1
        state["step"] += 1
        inter = state["sym"][slot][:buf_rows]
    else:
        inter = state["inter"]
    if events:
        events[0].record()
    rowscan_device(d_band, w, rows, spread, inter[halo_top:halo_top + rows], stride=stride, offset=offset,
                   test_above=test_above)
    if events:
        events[1].record()
    if world > 1:
        if state["halo"] == "peer":
            pull_halos(state, slot, rank)
        else:
            exchange_halos(inter, h, world, rank, spread, group)
    if events:
        events[2].record()
    envelope_device(inter, w, buf_rows, halo_top, halo_top + rows, spread, asymmetric, out, workspace=state["ws"])
    if events:
        events[3].record()
    return out


# ---------------------------------------------------------------- glyph / atlas front-end
def atlas_pack(wh, atlas_w, pad=1):
    """Skyline bottom-left packing (chaqsdf_atlas_pack; host only). wh: [n, 2] widths and heights.
    Returns (xy [n, 2] uint32, atlas_h)."""
    import numpy as np
    wh = np.ascontiguousarray(wh, np.uint32).reshape(-1, 2)
    xy = np.zeros_like(wh)
    ah = C.c_size_t(0)
    _check(lib().chaqsdf_atlas_pack(len(wh), wh.ctypes.data_as(C.c_void_p), atlas_w, pad, xy.ctypes.data_as(C.c_void_p),
                                    C.byref(ah)))
    return xy, int(ah.value)


def glyph_boxes_device(d_glyphs, n, w, h, test_above=True, margin=0, stream=None):
    import torch
    boxes = torch.empty((n, 4), dtype=torch.int32, device=d_glyphs.device)
    _check(lib().chaqsdf_glyph_boxes_device(_ptr(d_glyphs), n, w, h, int(test_above), margin, _ptr(boxes), _stream_ptr(stream)))
    return boxes


def atlas_blit_device(d_sdf, n, w, h, d_boxes, d_xy, d_atlas, atlas_w, stream=None):
    _check(lib().chaqsdf_atlas_blit_device(_ptr(d_sdf), n, w, h, _ptr(d_boxes), _ptr(d_xy), _ptr(d_atlas), atlas_w,
                                           _stream_ptr(stream)))
    return d_atlas


def atlas_u8(glyphs, n, w, h, spread=16, asymmetric=False, test_above=True, atlas_w=4096, pad=1, rows_cap=None):
    """chaqsdf_atlas_u8: n host canvases [n, h, w] -> (atlas [atlas_h, atlas_w] uint8, boxes [n, 4], xy [n, 2])."""
    import numpy as np
    glyphs = np.ascontiguousarray(glyphs, np.uint8)
    cap = rows_cap if rows_cap is not None else max(1, (n * (w + pad) * (h + pad) + atlas_w - 1) // atlas_w + h + pad)
    atlas = np.empty((cap, atlas_w), np.uint8)
    boxes = np.empty((n, 4), np.int32)
    xy = np.empty((n, 2), np.uint32)
    ah = C.c_size_t(0)
    _check(lib().chaqsdf_atlas_u8(glyphs.ctypes.data_as(C.c_void_p), n, w, h, int(test_above), spread, int(asymmetric), atlas_w, pad,
                                  atlas.ctypes.data_as(C.c_void_p), cap, C.byref(ah), boxes.ctypes.data_as(C.c_void_p),
                                  xy.ctypes.data_as(C.c_void_p)))
    return atlas[:ah.value], boxes, xy


# ---------------------------------------------------------------- transposed row bands (SURVEY.md 8(e) "row bands throughout")
def transpose_u8_device(src_ptr, src_pitch, rows, cols, dst_ptr, dst_pitch, stream=None):
    """dst[c*dst_pitch + r] = src[r*src_pitch + c] on the device (raw device addresses)."""
    _check(lib().chaqsdf_transpose_u8_device(C.c_void_p(src_ptr), src_pitch, rows, cols, C.c_void_p(dst_ptr), dst_pitch,
                                             _stream_ptr(stream)))


def envelope_cols_device(d_inter, w, buf_rows, out_begin, out_end, col_begin, col_end, spread, asymmetric, d_out,
                         workspace=None, stream=None):
    nb = int(lib().chaqsdf_envelope_workspace_bytes(w, buf_rows, spread))
    if workspace is None and nb:
        workspace = _workspace(nb, d_inter.device)
    _check(lib().chaqsdf_envelope_cols_device(_ptr(d_inter), w, buf_rows, out_begin, out_end, col_begin, col_end, spread,
                                              int(asymmetric), _ptr(d_out), _ptr(workspace),
                                              workspace.numel() if workspace is not None else 0, _stream_ptr(stream)))
    return d_out


def sdf_u8_tband_device(d_ext, w, ext_rows, halo_top, rows, spread, asymmetric=False, test_above=True, bufs=None, out=None):
    """The transposed-band computation on one device. d_ext: [ext_rows][w] input bytes = `halo_top` context rows, the
    band's `rows` rows, then the context rows below (context = `spread` rows of the image, fewer at an image edge).
    Transposes, runs pass 1 along the original columns (length ext_rows) and pass 2 along the original rows (length w,
    whole inside the band) for the band's columns only, and transposes back. Returns [rows][w] bytes.
    bufs: dict that keeps the scratch tensors between calls."""
    import torch
    dev = d_ext.device
    bufs = {} if bufs is None else bufs
    key = (w, ext_rows, spread)
    if bufs.get("key") != key:
        bufs.clear()
        bufs["key"] = key
        bufs["xt"] = torch.empty((w, ext_rows), dtype=torch.uint8, device=dev)
        bufs["inter"] = torch.empty((w, inter_row_bytes(ext_rows, spread)), dtype=torch.uint8, device=dev)
        bufs["out_t"] = torch.empty((w, ext_rows), dtype=torch.uint8, device=dev)
        bufs["ws"] = _workspace(int(lib().chaqsdf_envelope_workspace_bytes(ext_rows, w, spread)), dev)
    if out is None:
        out = torch.empty((rows, w), dtype=torch.uint8, device=dev)
    xt, inter, out_t = bufs["xt"], bufs["inter"], bufs["out_t"]
    transpose_u8_device(d_ext.data_ptr(), w, ext_rows, w, xt.data_ptr(), ext_rows)
    rowscan_device(xt, ext_rows, w, spread, inter, test_above=test_above)
    envelope_cols_device(inter, ext_rows, w, 0, w, halo_top, halo_top + rows, spread, asymmetric, out_t, workspace=bufs["ws"])
    transpose_u8_device(out_t.data_ptr() + halo_top, ext_rows, w, rows, out.data_ptr(), w)
    return out


def tband_setup(state, w, h, spread, rank, world, group, dev, exchange=None):
    """Buffers of the transposed-band path: the INPUT bands live in symmetric (peer-mapped) allocations, two slots
    alternating by call, laid out like the halo path's intermediates (row 0 = image row r0 - halo_top), so that a rank
    pulls its context rows of the input straight out of its neighbours' buffers."""
    import torch
    key = ("tband", w, h, spread, rank, world)
    if state.get("key") == key:
        return state
    state.clear()
    state["key"] = key
    plans = [halo_plan(h, world, r, spread) for r in range(world)]
    r0, r1, halo_top, halo_bot, _, _ = plans[rank]
    max_rows = max(p[2] + (p[1] - p[0]) + p[3] for p in plans)
    mode = exchange or os.environ.get("CHAQSDF_HALO") or ("peer" if dev.type == "cuda" and world > 1 else "p2p")
    if mode == "peer" and world > 1:
        ok = 1
        try:
            import torch.distributed as dist
            import torch.distributed._symmetric_memory as symm
            sym = symm.empty((2, max_rows, w), dtype=torch.uint8, device=dev)
            hdl = symm.rendezvous(sym, group if group is not None else dist.group.WORLD)
            bufs = {peer: hdl.get_buffer(peer, (2, max_rows, w), torch.uint8) for (peer, _, _) in plans[rank][4]}
            state.update(sym=sym, hdl=hdl, peer_bufs=bufs)
        except Exception as e:
            ok = 0
            state["peer_error"] = repr(e)
        import torch.distributed as dist
        flag = torch.tensor([ok], dtype=torch.int32, device=dev)
        dist.all_reduce(flag, op=dist.ReduceOp.MIN, group=group)
        if not bool(flag.item()):
            mode = "p2p"
    else:
        mode = "p2p"
    if mode != "peer":
        state["sym"] = torch.empty((2, max_rows, w), dtype=torch.uint8, device=dev)
    state.update(plans=plans, step=0, halo=mode, resident=False, bufs={},
                 out=torch.empty((r1 - r0, w), dtype=torch.uint8, device=dev))
    return state


def tband_input_view(state, rank, slot=None):
    """Where this rank's band of the input goes for the next call (or for slot `slot`): a [rows][w] view."""
    r0, r1, halo_top, _, _, _ = state["plans"][rank]
    s = (state["step"] & 1) if slot is None else slot
    return state["sym"][s][halo_top:halo_top + (r1 - r0)]


def sdf_u8_sharded_t(d_band, w, h, spread, asymmetric=False, rank=0, world=1, group=None, state=None, test_above=True,
                     events=None, exchange=None):
    """One image of h rows sharded by row bands over `world` ranks, transposed-band variant: every rank pulls `spread`
    rows of the INPUT from its neighbours (peer-mapped buffers over NVLink after a pairwise device-side signal, or
    NCCL / gloo point to point), then works alone: pass 1 runs along the original columns of band + context, pass 2
    along the original rows, which are whole inside the band, so the expensive pass does exactly 1/world of the work
    and waits for nobody's kernels. d_band: this rank's rows, or None when the band was written into
    tband_input_view(state, rank) (state from tband_setup). Returns this rank's output rows."""
    if state is None:
        state = {}
    dev = d_band.device if d_band is not None else state["sym"].device
    tband_setup(state, w, h, spread, rank, world, group, dev, exchange)
    r0, r1, halo_top, halo_bot, _, _ = state["plans"][rank]
    rows = r1 - r0
    ext_rows = halo_top + rows + halo_bot
    slot = state["step"] & 1
    state["step"] += 1
    ext = state["sym"][slot][:ext_rows]
    if d_band is not None and not state["resident"]:
        ext[halo_top:halo_top + rows].copy_(d_band, non_blocking=True)
    if events:
        events[0].record()
    if world > 1:
        if state["halo"] == "peer":
            pull_halos(state, slot, rank)
        else:
            exchange_halos(ext, h, world, rank, spread, group)
    if events:
        events[1].record()
    import torch  # noqa: F401
    bufs = state["bufs"]
    # the four launches of sdf_u8_tband_device, with an event between the scan and the envelope side
    key = (w, ext_rows, spread)
    if bufs.get("key") != key:
        sdf_u8_tband_device(ext, w, ext_rows, halo_top, rows, spread, asymmetric, test_above, bufs, state["out"])  # allocates
        if events:
            events[2].record()
            events[3].record()
        return state["out"]
    xt, inter, out_t = bufs["xt"], bufs["inter"], bufs["out_t"]
    transpose_u8_device(ext.data_ptr(), w, ext_rows, w, xt.data_ptr(), ext_rows)
    rowscan_device(xt, ext_rows, w, spread, inter, test_above=test_above)
    if events:
        events[2].record()
    envelope_cols_device(inter, ext_rows, w, 0, w, halo_top, halo_top + rows, spread, asymmetric, out_t, workspace=bufs["ws"])
    transpose_u8_device(out_t.data_ptr() + halo_top, ext_rows, w, rows, state["out"].data_ptr(), w)
    if events:
        events[3].record()
    return state["out"]


# ---------------------------------------------------------------- exact fields over several GPUs (all-to-all, variant E1)
_FIELD_DT = None


def col_part(w, nparts, q):
    """Columns [c0, c1) of part q when a w-column image is cut into nparts column parts (chaqsdf_col_part)."""
    c0, c1 = C.c_size_t(), C.c_size_t()
    lib().chaqsdf_col_part(w, nparts, q, C.byref(c0), C.byref(c1))
    return c0.value, c1.value


def rowscan_scatter_device(d_img, w, rows, row0, part_ptrs, stride=1, offset=0, test_above=True, stream=None):
    """Pass 1 of the exact path on `rows` rows starting at image row row0, stored by column part straight into the
    buffers at part_ptrs (device addresses, one per part; peers' buffers when they are peer-mapped)."""
    arr = (C.c_void_p * len(part_ptrs))(*part_ptrs)
    _check(lib().chaqsdf_rowscan_scatter_device(_ptr(d_img), w, rows, stride, offset, int(test_above), row0, len(part_ptrs),
                                                arr, _stream_ptr(stream)))


def fields_from_inter_device(d_inter, w, h, spread=64, asymmetric=False, want=("d2_in", "d2_out", "out"), stream=None):
    """Pass 2 of the exact path on a [h][w] uint16 intermediate (an image, or one column part of one)."""
    import torch
    dev = d_inter.device
    dt = {"d2_in": torch.int32, "d2_out": torch.int32, "f_in": torch.float32, "f_out": torch.float32,
          "sdf": torch.float32, "out": torch.uint8}
    res = {k: torch.empty((h, w), dtype=dt[k], device=dev) for k in want}
    ws = _workspace(int(lib().chaqsdf_fields_from_inter_workspace_bytes(w, h)), dev)
    g = lambda k: _ptr(res.get(k))  # noqa: E731
    _check(lib().chaqsdf_fields_from_inter_device(_ptr(d_inter), w, h, spread, int(asymmetric), g("d2_in"), g("d2_out"),
                                                  g("f_in"), g("f_out"), g("sdf"), g("out"), _ptr(ws), ws.numel(),
                                                  _stream_ptr(stream)))
    return res


def sdf_fields_a2a(d_band, w, h, spread=64, asymmetric=False, rank=0, world=1, group=None, state=None, stride=1, offset=0,
                   test_above=True, want=("d2_in", "d2_out", "out"), events=None, mode=None):
    """The exact (unwindowed) fields of one image over `world` ranks, one process per GPU: row bands for pass 1, an
    all-to-all to column parts, pass 2 on whole columns (SURVEY.md 8(e) variant E1). d_band holds this rank's rows
    (shard_rows); the result holds ALL rows of this rank's columns col_part(w, world, rank).
    mode "store" (default on GPUs): pass 1 writes its rows straight into every rank's peer-mapped part buffer over
    NVLink -- the all-to-all is the kernel's store -- then a pairwise device-side signal; mode "a2a": pass 1 into a
    local buffer, packed blocks through torch.distributed (NCCL all_to_all_single; also the gloo-free fallback).
    `state` (a dict) keeps the buffers and the peer mapping between calls; state["a2a"] says which mode runs."""
    import torch
    import torch.distributed as dist
    dev = d_band.device
    r0, r1 = shard_rows(h, world, rank)
    rows = r1 - r0
    c0, c1 = col_part(w, world, rank)
    pw = c1 - c0
    if state is None:
        state = {}
    key = ("a2a", w, h, rank, world)
    if state.get("key") != key:
        state.clear()
        state["key"] = key
        m = mode or os.environ.get("CHAQSDF_A2A") or ("store" if dev.type == "cuda" and world > 1 else "a2a")
        maxpw = max(col_part(w, world, q)[1] - col_part(w, world, q)[0] for q in range(world))
        if m == "store":
            ok = 1
            try:
                import torch.distributed._symmetric_memory as symm
                sym = symm.empty((2, h * maxpw), dtype=torch.int16, device=dev)
                hdl = symm.rendezvous(sym, group if group is not None else dist.group.WORLD)
                ptrs = [[hdl.get_buffer(q, (2, h * maxpw), torch.int16)[s].data_ptr() for q in range(world)] for s in range(2)]
                state.update(sym=sym, hdl=hdl, ptrs=ptrs, step=0)
            except Exception as e:  # noqa: BLE001 -- every rank falls back together
                ok = 0
                state["peer_error"] = repr(e)
            flag = torch.tensor([ok], dtype=torch.int32, device=dev)
            dist.all_reduce(flag, op=dist.ReduceOp.MIN, group=group)
            if not bool(flag.item()):
                m = "a2a"
        if m != "store":
            state["part"] = torch.empty((h * maxpw,), dtype=torch.int16, device=dev)
            state["local"] = torch.empty((rows, w), dtype=torch.int16, device=dev)
        state["a2a"] = m
    if events:
        events[0].record()
    if state["a2a"] == "store":
        slot = state["step"] & 1
        state["step"] += 1
        rowscan_scatter_device(d_band, w, rows, r0, state["ptrs"][slot], stride=stride, offset=offset, test_above=test_above)
        if events:
            events[1].record()
        hdl = state["hdl"]
        for q in range(world):
            if q != rank:
                hdl.put_signal(q, slot, 20000)
        for q in range(world):
            if q != rank:
                hdl.wait_signal(q, slot, 20000)
        part = state["sym"][slot][:h * pw].view(h, pw)
    else:
        local = state["local"]
        rowscan_scatter_device(d_band, w, rows, 0, [local.data_ptr()], stride=stride, offset=offset, test_above=test_above)
        if events:
            events[1].record()
        part = state["part"][:h * pw].view(h, pw)
        if world == 1:
            part.copy_(local)
        else:
            # (as bytes: NCCL has no 16-bit integer type)
            send = [local[:, a:b].contiguous().view(torch.uint8) for (a, b) in (col_part(w, world, q) for q in range(world))]
            recv = []
            for q in range(world):
                q0, q1 = shard_rows(h, world, q)
                recv.append(part[q0:q1].view(torch.uint8))  # rows of rank q, my columns: contiguous in the part buffer
            dist.all_to_all(recv, send, group=group)
    if events:
        events[2].record()
    if pw == 0:  # more ranks than 32-column parts: nothing to do here
        dtm = {"d2_in": torch.int32, "d2_out": torch.int32, "f_in": torch.float32, "f_out": torch.float32,
               "sdf": torch.float32, "out": torch.uint8}
        return {k: torch.empty((h, 0), dtype=dtm[k], device=dev) for k in want}
    res = fields_from_inter_device(part, pw, h, spread=spread, asymmetric=asymmetric, want=want)
    if events:
        events[3].record()
    return res
