"""Developer probe: host-fed commit (nlc_commit_from_values) with pinned vs pageable host columns."""
import ctypes as C, os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import __graft_entry__ as graft
from bench import synth_values
import torch
pkg = graft.load_package(); P2, ffi = pkg.plonky2, pkg._ffi; L = ffi.lib()
ctx = P2.Context(0)
for (log_n, W, r) in [(16, 135, 3), (16, 418, 1), (20, 400, 1)]:
    n = 1 << log_n
    vals = synth_values(W, n, 1)
    pinned = torch.empty((W, n), dtype=torch.int64).pin_memory(); pinned.numpy().view(np.uint64)[:] = vals
    cap = np.zeros((16, 4), np.uint64)
    for name, base in (("pinned", pinned.numpy().view(np.uint64)), ("pageable", vals)):
        tab = (ffi.u64p * W)(*[base[j].ctypes.data_as(ffi.u64p) for j in range(W)])
        ts = []
        for it in range(6):
            h = ffi.vp(); t0 = time.perf_counter()
            ctx._check(L.nlc_commit_from_values(ctx._h, tab, W, log_n, r, 4, C.byref(h)))
            ctx._check(L.nlc_batch_cap(h, cap.ctypes.data_as(ffi.u64p)))
            ts.append(time.perf_counter() - t0); L.nlc_batch_free(h)
        print(f"2^{log_n}x{W} r{r} {name}: {np.median(ts[2:])*1e3:.2f} ms (H2D {8*W*n/1e6:.0f} MB)", flush=True)
# D2H side: materialising the struct fields of a PolynomialBatch (coefficients, row-major leaves)
log_n, W, r = 16, 135, 3
n, N = 1 << log_n, 1 << (log_n + r)
vals = synth_values(W, n, 1)
tab = (ffi.u64p * W)(*[vals[j].ctypes.data_as(ffi.u64p) for j in range(W)])
h = ffi.vp(); ctx._check(L.nlc_commit_from_values(ctx._h, tab, W, log_n, r, 4, C.byref(h)))
for name, mk in (("pinned", lambda shape: torch.empty(shape, dtype=torch.int64).pin_memory().numpy().view(np.uint64)),
                 ("pageable", lambda shape: np.zeros(shape, np.uint64))):
    co = mk((W, n)); lv = mk((N, W))
    ctab = (ffi.u64p * W)(*[co[j].ctypes.data_as(ffi.u64p) for j in range(W)])
    for fn, args, mb in (("nlc_batch_coeffs", (ctab,), 8 * W * n / 1e6), ("nlc_batch_leaves", (lv.ctypes.data_as(ffi.u64p),), 8 * W * N / 1e6)):
        ts = []
        for it in range(4):
            t0 = time.perf_counter(); ctx._check(getattr(L, fn)(h, *args)); ts.append(time.perf_counter() - t0)
        print(f"{fn} {name}: {np.median(ts[1:])*1e3:.2f} ms ({mb:.0f} MB, {mb/np.median(ts[1:])/1e3:.1f} GB/s)", flush=True)
    if name == "pinned": ref_co, ref_lv = co.copy(), lv.copy()
    else: print("pageable == pinned:", np.array_equal(co, ref_co), np.array_equal(lv, ref_lv))
