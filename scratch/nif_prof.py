import os, sys; sys.path.insert(0, '.')
os.environ.setdefault("B200PT_LIB", os.path.abspath("ipu_path_trace_b200/libb200pt_prof.so"))
import ctypes as C, numpy as np, torch
import ipu_path_trace_b200 as pt
n = int(sys.argv[1]) if len(sys.argv) > 1 else 1 << 20
t = pt.PathTracer(64, 64); t.init_nif_weights(pt.synthetic_model())
u = torch.rand(n, device="cuda"); v = torch.rand(n, device="cuda"); out = torch.empty(n * 3, device="cuda")
t.nif_infer_device(n, u.data_ptr(), v.data_ptr(), out.data_ptr(), pt.NIF_TCGEN05, 2)
ms = t.nif_infer_device(n, u.data_ptr(), v.data_ptr(), out.data_ptr(), pt.NIF_TCGEN05, 1)
buf = (C.c_ulonglong * (148 * 16))()
assert t.lib.b200pt_debug_tc_profile(buf, 148 * 16) == 0
a = np.array(buf, dtype=np.float64).reshape(148, 16)
sel = a[:, 6] > 0; m = a[sel].mean(0); tiles = m[6]
print(f"ms {ms:.3f} tiles/CTA {tiles:.1f}")
print("MMA thread  per tile cycles: total %.0f  wait feat %.0f  acc_empty %.0f  act_ready %.0f  w_full %.0f  out_empty %.0f  (issue+other %.0f)" % (
    m[0]/tiles, m[1]/tiles, m[2]/tiles, m[3]/tiles, m[4]/tiles, m[5]/tiles, (m[0]-m[1:6].sum())/tiles))
print("  peer w_full wait %.0f ; producer (all CTAs) wait w_empty %.0f of %.0f per tile" % (m[7]/tiles, a[:,13].mean()/tiles, a[:,14].mean()/tiles))
print("  MMA+commit issue block %.0f per tile" % (m[15]/tiles))
print("epilogue w0 per tile cycles: total %.0f  wait acc_full %.0f  work %.0f (per job %.0f)  out wait %.0f  out work %.0f" % (
    m[8]/tiles, m[9]/tiles, m[10]/tiles, m[10]/tiles/12, m[11]/tiles, m[12]/tiles))
