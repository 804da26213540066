"""Timeline of one 128-row tile of nif_tc_kernel (profile build: make -C ipu_path_trace_b200/csrc prof): when each weight chunk's MMAs were
issued by which warp, how long it waited for what, and when the epilogue saw / released each accumulator."""
import os, sys; sys.path.insert(0, '.')
os.environ.setdefault("B200PT_LIB", os.path.abspath("ipu_path_trace_b200/libb200pt_prof.so"))
import ctypes as C, numpy as np, torch
import ipu_path_trace_b200 as pt
n = 1 << 22
t = pt.PathTracer(64, 64); t.init_nif_weights(pt.synthetic_model())
u = torch.rand(n, device="cuda"); v = torch.rand(n, device="cuda"); out = torch.empty(n * 3, device="cuda")
t.nif_infer_device(n, u.data_ptr(), v.data_ptr(), out.data_ptr(), pt.NIF_TCGEN05, 2)
t.nif_infer_device(n, u.data_ptr(), v.data_ptr(), out.data_ptr(), pt.NIF_TCGEN05, 1)
buf = (C.c_ulonglong * 256)()
assert t.lib.b200pt_debug_tc_trace(buf, 256) == 0
a = np.array(buf, dtype=np.int64)
iss = a[:128].reshape(32, 4); epi = a[128:128 + 32].reshape(16, 2)
nch = int((iss[:, 0] > 0).sum())
t0 = iss[0, 0]
print("chunk | loop top | barriers passed (+wait) | turn taken (+wait) | issued (+issue) | gap to previous chunk's issue end")
prev_end = None
for c in range(nch):
    r = iss[c] - t0
    gap = "" if prev_end is None else f"{r[2] - prev_end:6d}"
    print(f"{c:3d} | {r[0]:7d} | {r[1]:7d} (+{r[1]-r[0]:5d}) | {r[2]:7d} (+{r[2]-r[1]:5d}) | {r[3]:7d} (+{r[3]-r[2]:5d}) | {gap}")
    prev_end = r[3]
print("tile total (first chunk top -> last chunk issued):", iss[nch - 1, 3] - t0)
print("epilogue warp 0: job | accumulator full seen | hand-over done (+work)")
for j in range(12):
    if epi[j, 0] > 0: print(f"{j:3d} | {epi[j,0]-t0:7d} | {epi[j,1]-t0:7d} (+{epi[j,1]-epi[j,0]:5d})")
