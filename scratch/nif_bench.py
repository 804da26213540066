"""C4 microbenchmark: NIF MLP on n (default 1M) escaped-ray directions resident in HBM."""
import sys; sys.path.insert(0, '.')
import json, numpy as np, torch
import ipu_path_trace_b200 as pt
n = int(sys.argv[1]) if len(sys.argv) > 1 else 1 << 20
reps = int(sys.argv[2]) if len(sys.argv) > 2 else 20
impl = {"tc": pt.NIF_TCGEN05, "simt": pt.NIF_SIMT}[sys.argv[3] if len(sys.argv) > 3 else "tc"]
t = pt.PathTracer(64, 64); m = pt.synthetic_model(); t.init_nif_weights(m)
g = torch.Generator(device="cuda"); g.manual_seed(1234)
u = torch.rand(n, device="cuda", generator=g); v = torch.rand(n, device="cuda", generator=g)
out = torch.empty(n * 3, device="cuda")
torch.cuda.synchronize()
t.nif_infer_device(n, u.data_ptr(), v.data_ptr(), out.data_ptr(), impl, 3)
ms = t.nif_infer_device(n, u.data_ptr(), v.data_ptr(), out.data_ptr(), impl, reps) / reps
fl = t.nif_flops_per_eval()
print(json.dumps({"n": n, "ms": ms, "rows_per_s": n / ms * 1e3, "tflops": n * fl / ms / 1e9, "finite": bool(torch.isfinite(out).all())}))
