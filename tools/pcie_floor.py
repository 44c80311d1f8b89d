"""What the host link allows: H2D and D2H of one 16384^2 byte image each, chunked, on two streams at once.
The e2e number of chaqsdf_u8 cannot beat this. usage: python tools/pcie_floor.py"""
import torch
n = 16384 * 16384
h_in = torch.empty(n, dtype=torch.uint8).pin_memory()
h_out = torch.empty(n, dtype=torch.uint8).pin_memory()
d_in = torch.empty(n, dtype=torch.uint8, device="cuda")
d_out = torch.empty(n, dtype=torch.uint8, device="cuda")
s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()
def run(chunk, both):
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    a.record()
    s1.wait_event(a); s2.wait_event(a)
    for o in range(0, n, chunk):
        with torch.cuda.stream(s1):
            d_in[o:o + chunk].copy_(h_in[o:o + chunk], non_blocking=True)
        if both:
            with torch.cuda.stream(s2):
                h_out[o:o + chunk].copy_(d_out[o:o + chunk], non_blocking=True)
    torch.cuda.current_stream().wait_stream(s1); torch.cuda.current_stream().wait_stream(s2)
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b)
for chunk in (n, 32 << 20, 8 << 20):
    for both in (False, True):
        run(chunk, both)
        t = min(run(chunk, both) for _ in range(5))
        print(f"chunk {chunk >> 20} MiB  {'H2D + D2H concurrently' if both else 'H2D alone'}: {t:.3f} ms  ({n / t / 1e6:.1f} GB/s per direction)")
