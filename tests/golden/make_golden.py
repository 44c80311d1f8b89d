"""Regenerates tests/golden/*.npz from the reference tree. Run in the build container only
(needs /root/reference and oracle/_ref/libchaqref.so built by oracle/Makefile).

  sample_pair.npz : the reference's single golden vector (README.md:8): image/sample_input.png decoded as
                    the CLI decodes it (stbi_load req_comp=2, sdfgen.c:253-255) and image/sample_output.png's
                    pixels; flags -s 100 -al.
  ref_vectors.npz : outputs of the stock OpenMP build (ref_sdf_u8 = sdfgen.c:264-300) on small seeded
                    inputs, for flag/threshold/stride coverage: a 96x96 (Y,A) gradient image with
                    126/127/128 patches x flags {default,-l,-n,-a,-ln,-aln} x spread {7,64}; plus D2/float
                    fields of a 64x64 noise mask.
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(ROOT, "oracle"))
from oracle import Ref  # noqa: E402

HERE = os.path.dirname(os.path.abspath(__file__))


def gradient_ya(n=96, seed=7):
    rng = np.random.default_rng(seed)
    y, x = np.mgrid[0:n, 0:n]
    lum = ((x * 255) // (n - 1)).astype(np.uint8)
    alp = ((y * 255) // (n - 1)).astype(np.uint8)
    ya = np.stack([lum, alp], -1).copy()
    for k, v in enumerate((126, 127, 128)):
        ya[10 + 20 * k:22 + 20 * k, 8:20, 0] = v
        ya[10 + 20 * k:22 + 20 * k, 70:82, 1] = v
    blobs = rng.integers(0, n - 8, (12, 2))
    for (by, bx) in blobs:
        ya[by:by + 6, bx:bx + 5, :] = rng.integers(0, 256, 2, dtype=np.uint8)
    return ya


def main():
    from PIL import Image
    r = Ref()
    ya, ncomp = r.load_ya("/root/reference/image/sample_input.png")
    gold = np.array(Image.open("/root/reference/image/sample_output.png"))
    assert gold.dtype == np.uint8 and gold.shape == (200, 200)
    np.savez_compressed(os.path.join(HERE, "sample_pair.npz"), ya=ya, out=gold, spread=100, asymmetric=1,
                        offset=0, test_above=1)

    g = gradient_ya()
    n = g.shape[0]
    vec = {"grad_ya": g}
    flags = {"default": (1, True, False), "l": (0, True, False), "n": (1, False, False), "a": (1, True, True),
             "ln": (0, False, False), "aln": (0, False, True)}
    for name, (off, above, asym) in flags.items():
        for s in (7, 64):
            vec[f"grad_{name}_s{s}"] = r.sdf_u8(g, n, n, 2, off, above, s, asym)
    rng = np.random.default_rng(3)
    m = ((rng.random((64, 64)) < 1 / 40) * 255).astype(np.uint8)
    out, sdf, din, dout = r.sdf_fields(m, 64, 64, 1, 0, True, 16, False)
    vec.update(noise_img=m, noise_out=out, noise_sdf=sdf, noise_din=din, noise_dout=dout)
    np.savez_compressed(os.path.join(HERE, "ref_vectors.npz"), **vec)
    print("wrote", os.listdir(HERE))


if __name__ == "__main__":
    main()
