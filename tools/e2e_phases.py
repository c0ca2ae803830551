import ctypes as C, os, sys, time
import numpy as np
sys.path.insert(0, '/root/repo')
import __graft_entry__ as graft
from bench import synth_values
import torch
pkg = graft.load_package(); P2, ffi = pkg.plonky2, pkg._ffi; L = ffi.lib()
ctx = P2.Context(0)
log_n, W, r = 16, 135, 3
n = 1 << log_n
vals = synth_values(W, n, 1)
pinned = torch.empty((W, n), dtype=torch.int64).pin_memory(); pinned.numpy().view(np.uint64)[:] = vals
base = pinned.numpy().view(np.uint64)
tab = (ffi.u64p * W)(*[base[j].ctypes.data_as(ffi.u64p) for j in range(W)])
dev = pinned.cuda()
cap = np.zeros((16, 4), np.uint64)
def e2e():
    h = ffi.vp(); t0 = time.perf_counter()
    ctx._check(L.nlc_commit_from_values(ctx._h, tab, W, log_n, r, 4, C.byref(h)))
    t1 = time.perf_counter()
    ctx._check(L.nlc_batch_cap(h, cap.ctypes.data_as(ffi.u64p)))
    t2 = time.perf_counter(); L.nlc_batch_free(h); return t1 - t0, t2 - t1
def res():
    h = ffi.vp(); t0 = time.perf_counter()
    ctx._check(L.nlc_commit_from_values_dev(ctx._h, ffi.vp(dev.data_ptr()), W, log_n, r, 4, C.byref(h)))
    ctx.synchronize(); t1 = time.perf_counter(); L.nlc_batch_free(h); return t1 - t0
for _ in range(3): e2e(); res()
print("e2e commit / cap ms:", np.median([e2e() for _ in range(10)], axis=0) * 1e3)
print("resident wall ms:", np.median([res() for _ in range(10)]) * 1e3)
ctx.timing_enable(True); ctx.timing_read()
for _ in range(5): e2e()
print("e2e phases:", {k: v / 5 for k, v in ctx.timing_read().items()})
for _ in range(5): res()
print("resident phases:", {k: v / 5 for k, v in ctx.timing_read().items()})
