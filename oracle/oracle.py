"""ctypes loader for the TEST-ONLY CPU checkers (see oracle/edt_oracle.c header).

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs import this.
Nothing under chaq-sdfgen_b200/ may.

  Oracle  -> oracle/liboracle.so      exact integer restatement (this repo's own code)
  Ref     -> oracle/_ref/libchaqref.so the unmodified reference hot path, compiled from the reference tree
             (oracle/Makefile); prebuilt copy travels to the GPU box, where /root/reference is absent.
"""
import ctypes as C
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
_u8p = np.ctypeslib.ndpointer(np.uint8, flags="C_CONTIGUOUS")
_i32p = np.ctypeslib.ndpointer(np.int32, flags="C_CONTIGUOUS")
_f32p = np.ctypeslib.ndpointer(np.float32, flags="C_CONTIGUOUS")
INF_D2 = np.int32(2**31 - 1)


def build(force=False):
    """(Re)build liboracle.so and, when the reference tree is present, _ref/libchaqref.so."""
    need = force or not os.path.exists(os.path.join(HERE, "liboracle.so"))
    src = os.path.join(HERE, "edt_oracle.c")
    lib = os.path.join(HERE, "liboracle.so")
    if not need and os.path.getmtime(src) > os.path.getmtime(lib):
        need = True
    if need or (os.path.exists("/root/reference/openmp/df.c")
                and not (os.path.exists(os.path.join(HERE, "_ref", "libchaqref.so"))
                         and os.path.exists(os.path.join(HERE, "_ref", "libchaqref_dbl.so")))):
        subprocess.run(["make", "-C", HERE, "-s"], check=True, stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)


def _opt(a, dtype):
    if a is None:
        return None
    assert a.dtype == dtype and a.flags.c_contiguous
    return a.ctypes.data_as(C.c_void_p)


class Oracle:
    def __init__(self):
        build()
        L = C.CDLL(os.path.join(HERE, "liboracle.so"))
        L.oracle_img_to_mask.argtypes = [_u8p, _u8p, C.c_size_t, C.c_size_t, C.c_size_t, C.c_size_t, C.c_int]
        L.oracle_edt_d2.argtypes = [_u8p, C.c_size_t, C.c_size_t, C.c_int, _i32p]
        L.oracle_edt_d2_bruteforce.argtypes = [_u8p, C.c_size_t, C.c_size_t, C.c_int, _i32p]
        L.oracle_sdf_fields.argtypes = [_u8p, C.c_size_t, C.c_size_t, C.c_size_t, C.c_size_t, C.c_int, C.c_size_t,
                                        C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
        L.oracle_sdf_fields.restype = C.c_int
        L.oracle_dist_transform_2d_binary.argtypes = [_f32p, C.c_size_t, C.c_size_t]
        L.oracle_dist_transform_2d_binary.restype = C.c_int
        L.oracle_syn_discs.argtypes = [_u8p, C.c_size_t, C.c_size_t, C.c_uint64, C.c_int]
        L.oracle_syn_noise.argtypes = [_u8p, C.c_size_t, C.c_size_t, C.c_uint64]
        L.oracle_syn_glyph.argtypes = [_u8p, C.c_size_t, C.c_uint64]
        L.oracle_signed_from_d2.argtypes = [C.c_int32, C.c_int32]
        L.oracle_signed_from_d2.restype = C.c_float
        L.oracle_byte_from_signed.argtypes = [C.c_float, C.c_size_t, C.c_int]
        L.oracle_byte_from_signed.restype = C.c_uint8
        self.L = L

    def mask(self, img, w, h, stride=1, offset=0, test_above=True):
        m = np.empty((h, w), np.uint8)
        self.L.oracle_img_to_mask(np.ascontiguousarray(img).reshape(-1), m.reshape(-1), w, h, stride, offset,
                                  int(test_above))
        return m

    def edt_d2(self, mask, seed_value=1, brute=False):
        h, w = mask.shape
        d2 = np.empty((h, w), np.int32)
        f = self.L.oracle_edt_d2_bruteforce if brute else self.L.oracle_edt_d2
        f(np.ascontiguousarray(mask, np.uint8).reshape(-1), w, h, int(seed_value), d2.reshape(-1))
        return d2

    def sdf(self, img, w, h, stride=1, offset=0, test_above=True, spread=64, asymmetric=False, fields=False):
        """Returns u8[h,w]; with fields=True returns (u8, sdf f32, d2_in i32, d2_out i32)."""
        img = np.ascontiguousarray(img, np.uint8).reshape(-1)
        out = np.empty((h, w), np.uint8)
        if not fields:
            rc = self.L.oracle_sdf_fields(img, w, h, stride, offset, int(test_above), spread, int(asymmetric),
                                          _opt(out, np.uint8), None, None, None)
            assert rc == 0
            return out
        sdf = np.empty((h, w), np.float32)
        din = np.empty((h, w), np.int32)
        dout = np.empty((h, w), np.int32)
        rc = self.L.oracle_sdf_fields(img, w, h, stride, offset, int(test_above), spread, int(asymmetric),
                                      _opt(out, np.uint8), _opt(sdf, np.float32), _opt(din, np.int32),
                                      _opt(dout, np.int32))
        assert rc == 0
        return out, sdf, din, dout

    def dist_transform_2d_binary(self, img_f32):
        h, w = img_f32.shape
        a = np.ascontiguousarray(img_f32, np.float32).copy()
        rc = self.L.oracle_dist_transform_2d_binary(a.reshape(-1), w, h)
        assert rc == 0
        return a

    def syn_discs(self, w, h=None, seed=0, n_discs=200):
        h = w if h is None else h
        img = np.empty((h, w), np.uint8)
        self.L.oracle_syn_discs(img.reshape(-1), w, h, seed, n_discs)
        return img

    def syn_glyphs(self, count, n=256, first=0):
        """SYN-GLYPH(first) .. SYN-GLYPH(first + count - 1) as one [count, n, n] array."""
        imgs = np.empty((count, n, n), np.uint8)
        for i in range(count):
            self.L.oracle_syn_glyph(imgs[i].reshape(-1), n, first + i)
        return imgs

    def syn_noise(self, w, h=None, seed=0):
        h = w if h is None else h
        img = np.empty((h, w), np.uint8)
        self.L.oracle_syn_noise(img.reshape(-1), w, h, seed)
        return img


class Ref:
    """The stock reference build. Square images only; exact only for side <= 4098."""

    def __init__(self, variant=""):
        """variant "": the unmodified build; "_dbl": parabola_intersect in double (oracle/Makefile), exact at any size."""
        build()
        path = os.path.join(HERE, "_ref", f"libchaqref{variant}.so")
        if not os.path.exists(path):
            raise FileNotFoundError(path)
        L = C.CDLL(path)
        L.ref_sdf_u8.argtypes = [_u8p, C.c_size_t, C.c_size_t, C.c_size_t, C.c_size_t, C.c_int, C.c_size_t, C.c_int,
                                 _u8p]
        L.ref_sdf_u8.restype = C.c_int
        L.ref_sdf_fields.argtypes = [_u8p, C.c_size_t, C.c_size_t, C.c_size_t, C.c_size_t, C.c_int, C.c_size_t,
                                     C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
        L.ref_sdf_fields.restype = C.c_int
        L.ref_dist_transform_2d.argtypes = [_f32p, C.c_size_t, C.c_size_t]
        L.ref_d2_field.argtypes = [_f32p, C.c_size_t, C.c_size_t]
        L.ref_load_ya.argtypes = [C.c_char_p, C.POINTER(C.c_int), C.POINTER(C.c_int), C.POINTER(C.c_int)]
        L.ref_load_ya.restype = C.c_void_p
        L.ref_free.argtypes = [C.c_void_p]
        L.ref_omp_max_threads.restype = C.c_int
        self.L = L

    @staticmethod
    def available():
        return os.path.exists(os.path.join(HERE, "_ref", "libchaqref.so")) or os.path.exists(
            "/root/reference/openmp/df.c")

    def sdf_u8(self, img, w, h, stride=1, offset=0, test_above=True, spread=64, asymmetric=False):
        assert w == h, "stock reference is undefined for non-square images (df.c:47, df.c:92)"
        out = np.empty((h, w), np.uint8)
        rc = self.L.ref_sdf_u8(np.ascontiguousarray(img, np.uint8).reshape(-1), w, h, stride, offset,
                               int(test_above), spread, int(asymmetric), out.reshape(-1))
        assert rc == 0
        return out

    def sdf_fields(self, img, w, h, stride=1, offset=0, test_above=True, spread=64, asymmetric=False):
        assert w == h
        out = np.empty((h, w), np.uint8)
        sdf = np.empty((h, w), np.float32)
        din = np.empty((h, w), np.float32)
        dout = np.empty((h, w), np.float32)
        rc = self.L.ref_sdf_fields(np.ascontiguousarray(img, np.uint8).reshape(-1), w, h, stride, offset,
                                   int(test_above), spread, int(asymmetric), _opt(out, np.uint8),
                                   _opt(sdf, np.float32), _opt(din, np.float32), _opt(dout, np.float32))
        assert rc == 0
        return out, sdf, din, dout

    def dist_transform_2d(self, img_f32):
        h, w = img_f32.shape
        assert w == h
        a = np.ascontiguousarray(img_f32, np.float32).copy()
        self.L.ref_dist_transform_2d(a.reshape(-1), w, h)
        return a

    def d2_field(self, img_f32):
        h, w = img_f32.shape
        assert w == h
        a = np.ascontiguousarray(img_f32, np.float32).copy()
        self.L.ref_d2_field(a.reshape(-1), w, h)
        return a

    def load_ya(self, path):
        w, h, n = C.c_int(), C.c_int(), C.c_int()
        p = self.L.ref_load_ya(path.encode(), C.byref(w), C.byref(h), C.byref(n))
        if not p:
            raise IOError(path)
        buf = np.ctypeslib.as_array(C.cast(p, C.POINTER(C.c_uint8)), shape=(h.value, w.value, 2)).copy()
        self.L.ref_free(p)
        return buf, n.value

    def omp_max_threads(self):
        return int(self.L.ref_omp_max_threads())
