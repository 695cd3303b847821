#!/usr/bin/env python3
"""Host <-> device copy rates of this box as awfl_time_step_host uses them: contiguous and member-slab strided
(cudaMemcpy2DAsync rows of 2/4/8/32 KB), one direction and both at once. Diagnostic for DESIGN.md 5 (e2e)."""
import ctypes as C
import json
import sys
import torch

rt = C.CDLL("libcudart.so.12")
H2D, D2H = 1, 2
rt.cudaMemcpy2DAsync.argtypes = [C.c_void_p, C.c_size_t, C.c_void_p, C.c_size_t, C.c_size_t, C.c_size_t, C.c_int, C.c_void_p]
rows, nens = 58 * 64, 4096
host = torch.empty(6, rows, nens, dtype=torch.float64).pin_memory()
host2 = torch.empty(6, rows, nens, dtype=torch.float64).pin_memory()
dev = torch.empty(6, rows, nens, dtype=torch.float64, device="cuda")
dev2 = torch.empty(6, rows, nens, dtype=torch.float64, device="cuda")
s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()
nbytes = host.numel() * 8


def copy2d(dst, src, kind, ne, st):
    for v in range(6):
        for e0 in range(0, nens, ne):
            rt.cudaMemcpy2DAsync(dst[v].data_ptr() + 8 * e0, 8 * nens, src[v].data_ptr() + 8 * e0, 8 * nens, 8 * ne, rows, kind, st.cuda_stream)


def timed(fn, reps=5):
    fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        fn()
        s1.synchronize(); s2.synchronize()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps


out = {}
for ne in ((1024,) if '--quick' in sys.argv else (4096, 1024, 512, 256)):   # --quick: 8 KB rows only (run one per GPU at once)
    out["h2d_row%dB" % (8 * ne)] = nbytes / timed(lambda: copy2d(dev, host, H2D, ne, s1)) / 1e6
    out["d2h_row%dB" % (8 * ne)] = nbytes / timed(lambda: copy2d(host2, dev2, D2H, ne, s2)) / 1e6
    out["both_row%dB_each" % (8 * ne)] = nbytes / timed(lambda: (copy2d(dev, host, H2D, ne, s1), copy2d(host2, dev2, D2H, ne, s2))) / 1e6
print(json.dumps(out))
