#!/usr/bin/env python
"""bench.py - path samples/s of the hot path on B200 (BASELINE.json metric, config C2/C3).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]

A "step" is one `path_trace` program of the reference's render loop (PathTracerApp.cpp:648-776):
samples-per-step (300) samples of every pixel of the 1104x1000 image through camera ray ->
path -> escaped-ray NIF -> accumulate. At N > 1 (torchrun, one rank per GPU) the SAME image is
dealt to the ranks as shuffled 16x16 pixel tiles (LoadBalancer re-read for GPUs), so total work
is fixed: "scaling": "strong". No collective runs inside the sample loop; one NCCL reduce of the
disjoint-support film per save interval is timed and reported separately.

`value`    whole-job samples/s with the worklist resident in HBM (path_trace + film accumulate).
`e2e`      the same through the reference-facing calls with HOST (page-locked) buffers every step, as the reference's
           double-buffered loop does it: path_trace, setup_next (H->D work list of the next step, during the render),
           read_results (D->H TraceRecords). Resident and e2e steps are timed in alternating blocks.
`config`   also carries BASELINE config C4 (`c4`), the README's full-length run through the C++ host (`cli_c2`) and, at
           8 GPUs, C5 (`cli_c5`); --no-extras skips them.
`roofline` the NIF tcgen05 kernel: algorithmic flops (NifModel.cpp:122-134 formula x escaped rays)
           / its CUDA-event time, against MEASURED_PEAKS.json bf16_tflops_sustained.
`cpu_baseline` the CPU oracle (oracle/, a port: the reference has no CPU NIF) on a bounded sample.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

W, H = 1104, 1000                 # README.md:82 of the reference
SAMPLES_PER_STEP = 300
SAVE_INTERVAL = 10
METRIC = "path samples/sec at 1104x1000 w/ NIF env light"
UNIT = "samples/s"


def measured_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return float(d["bf16_tflops_sustained"]), float(d["hbm_gbs"]), "measured (MEASURED_PEAKS.json, sustained)"
    return 1400.0, 6650.0, "fallback (B200_PROFILING.md)"


NIF_KERNEL_SOURCES = ("nif_tc.cu", "nif_common.cuh", "api_nif.cu")


def nif_source_hashes():
    import hashlib
    d = os.path.join(ROOT, "ipu_path_trace_b200", "csrc")
    return {f: hashlib.sha256(open(os.path.join(d, f), "rb").read()).hexdigest() for f in NIF_KERNEL_SOURCES}


def nif_traffic_bytes():
    """(bytes, note): dram__bytes_read.sum + dram__bytes_write.sum of ONE nif_tc_kernel launch of this very workload (one
    wavefront of C2) from the newest committed `ncu --set full` capture, profiles/r*_nif_tc_traffic.json - ncu cannot run
    inside the timed region. A capture is only used if it records the hashes of the kernel's sources and they are those of
    the files in this tree (scratch/make_traffic_json.py writes them); otherwise (None, why), never a stale constant."""
    import glob
    files = sorted(glob.glob(os.path.join(ROOT, "profiles", "r*_nif_tc_traffic.json")), reverse=True)
    if not files:
        return None, "no capture committed"
    d = json.load(open(files[0]))
    name = os.path.basename(files[0])
    if d.get("source_sha256") != nif_source_hashes():
        return None, f"profiles/{name} was captured from other kernel sources than this tree's (source_sha256 differs or is absent): re-capture"
    return float(d["dram_bytes_read_per_launch"]) + float(d["dram_bytes_write_per_launch"]), f"profiles/{name}"


class ClockSampler:
    """nvidia-smi clocks and throttle reasons DURING the timed region (B200_PROFILING.md)."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, device: int):
        self.device = device
        self.lines = []
        self.proc = None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "200",
                                          "-i", str(self.device)], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def stop(self) -> dict:
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        sm, smax, power, reasons = [], [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ln in self.lines:
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 9:
                continue
            try:
                sm.append(float(f[1])); smax.append(float(f[2])); power.append(float(f[3]))
            except ValueError:
                continue
            for nm, val in zip(names, f[5:9]):
                if val.lower().startswith("active"):
                    reasons.add(nm)
        if not sm:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["no samples"]}
        return {"sm_mhz": float(np.median(sm)), "sm_max_mhz": float(max(smax)), "power_w_max": float(max(power)),
                "samples": len(sm), "reasons": sorted(reasons)}


# ---------------------------------------------------------------------------------------------
# CPU legs (the only places bench.py executes oracle/)
# ---------------------------------------------------------------------------------------------
def workload_name(sps: int) -> str:
    """One string for both arms, so the driver sees the same `config.workload` on the GPU and the CPU line."""
    return f"C2: {W}x{H} spheres + NIF env light, samples-per-step {sps}"


def cpu_sample_run(target_seconds: float, seed: int = 1):
    """Times the CPU statement of the path on a bounded sample of the C2 workload: the whole 1104x1000 image for
    a few samples per pixel (a short calibration on a random subset of 16x16 pixel tiles sizes it to about
    `target_seconds`). Path tracing = the pinned restatement of the reference's codelets; NIF = the batched AVX
    GEMM statement (oracle/nif_fast.cpp, mode 2), all host threads. Returns (samples/s, description, threads)."""
    import ipu_path_trace_b200 as pt
    from oracle import oracle as O
    O.build(ref=False)
    orc = O.Oracle()
    cores = orc.use_all_host_threads()          # torch.distributed.run exports OMP_NUM_THREADS=1
    model = pt.synthetic_model()
    nif = O.NifLayers(model.embedding_dimension, model.max, model.mean, model.log_tone_map, model.layers)
    s = O.Settings(width=W, height=H, seed=seed)
    tiles = pt.shard_tiles(W, H, 4313)          # one 16x16 tile per shard (69 x 63 tiles)
    order = np.random.default_rng(7).permutation(len(tiles))
    cal = np.concatenate([tiles[i] for i in order[:512]])
    orc.render(cal[:4096], 0, 1, s, nif, mode=2)               # page in, spin the thread team up
    t0 = time.perf_counter()
    orc.render(cal, 0, 1, s, nif, mode=2)
    rate = len(cal) / (time.perf_counter() - t0)
    want = rate * target_seconds
    if want >= W * H:
        items = pt.create_work_list_for_image(W, H)
        spp = max(1, int(want / (W * H)))
        what = f"the whole {W}x{H} image x {spp} spp"
    else:
        n_tiles = max(512, int(want / 256))
        items = np.concatenate([tiles[i] for i in order[:n_tiles]])
        spp = 1
        what = f"{n_tiles} random 16x16 tiles of the {W}x{H} image ({len(items)} pixels) x 1 spp"
    t0 = time.perf_counter()
    _, esc, t_tr, t_nif = orc.render(items, 0, spp, s, nif, mode=2)
    dt = time.perf_counter() - t0
    n = len(items) * spp
    desc = (f"{what} = {n} path samples in {dt:.1f} s (trace {t_tr:.2f} s, NIF {t_nif:.2f} s = "
            f"{esc * 1.089283e-3 / max(t_nif, 1e-9):.0f} GFLOP/s fp32 batched GEMM), escaped {esc / n:.3f}")
    return n / dt, desc, cores


def run_reference(args) -> None:
    """--impl reference: the reference has no CPU implementation of the NIF and its Poplar IPUModel is
    unavailable here, so this arm times the CPU oracle: the reference's path-tracing algorithm (pinned
    bit-for-bit against the unmodified codelets, tests/test_oracle_golden.py) + the NIF restatement, on
    all host threads. Rank 0 only."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    vals = []
    desc, cores = "", 1
    per_step = args.cpu_seconds if args.cpu_seconds > 0 else max(2.0, min(20.0, 120.0 / max(1, args.steps + args.warmup)))
    for i in range(args.warmup + args.steps):
        v, desc, cores = cpu_sample_run(per_step)
        if i >= args.warmup:
            vals.append(v)
    value = float(np.mean(vals))
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": 1e3 * W * H * args.samples_per_step / value, "higher_is_better": True,
        "scaling": "strong", "vs_baseline": None, "dtype": "f32 (binary16 weights/activations, fp32 accumulate, AVX batched GEMM)",
        "data": "synthetic (built-in sphere scene; seeded synthetic NIF weights in the shipped topology)",
        "config": {"workload": workload_name(args.samples_per_step),
                   "note": "each step is a bounded sample of the workload; ms_per_step is extrapolated to a full step"},
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores, "kind": "port", "sample": desc},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


# ---------------------------------------------------------------------------------------------
# Extras the driver can see: BASELINE configs C4 and C5, and the reference's own run-level metric
# (the final `Samples/sec` line of PathTracerApp.cpp:782-791) through the C++ host with the README's command line.
# ---------------------------------------------------------------------------------------------
def c4_uv(n: int = 1 << 20, seed: int = 1234):
    """BASELINE config C4: n unit directions uniform on the sphere (Philox seed 1234) -> equirectangular (u,v)
    (codelets.cpp:333-347 in numpy; the microbenchmark only needs their distribution)."""
    g = np.random.Generator(np.random.Philox(seed))
    d = g.standard_normal((n, 3))
    d /= np.linalg.norm(d, axis=1, keepdims=True)
    u = np.arccos(np.clip(d[:, 1], -1, 1)) / np.pi
    phi = np.arctan2(d[:, 2], d[:, 0])
    phi[phi < 0] += 2 * np.pi
    return u.astype(np.float32), (phi / (2 * np.pi)).astype(np.float32)


def run_c4(tracer, flops_per_eval: int, reps: int = 20):
    import torch
    import ipu_path_trace_b200 as pt
    u, v = c4_uv()
    du, dv = torch.from_numpy(u).cuda(), torch.from_numpy(v).cuda()
    out = torch.empty(3 * len(u), dtype=torch.float32, device="cuda")
    torch.cuda.synchronize()
    tracer.nif_infer_device(len(u), du.data_ptr(), dv.data_ptr(), out.data_ptr(), pt.NIF_TCGEN05, 3)          # warm-up
    ms = tracer.nif_infer_device(len(u), du.data_ptr(), dv.data_ptr(), out.data_ptr(), pt.NIF_TCGEN05, reps) / reps
    burst = None
    pk = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(pk):
        burst = float(json.load(open(pk))["bf16_tflops"])
    tf = len(u) * flops_per_eval / (ms * 1e-3) / 1e12
    return {"workload": "C4: 1048576 escaped-ray directions (Philox 1234) through the 12-dim embedding + 6x320 fp16 MLP, resident in HBM",
            "reps": reps, "ms_per_batch": ms, "rays_per_s": len(u) / (ms * 1e-3), "tflops": tf,
            "frac_of_burst_peak": (tf / burst) if burst else None, "burst_peak_tflops": burst, "finite": bool(torch.isfinite(out).all())}


def run_cli(n_gpus: int, width: int, height: int, spp: int, sps: int, save_interval: int, tag: str, timeout: float):
    """host/b200_path_trace with the reference's command line (README.md:82); parses its log like a user would."""
    import re
    import tempfile
    import ipu_path_trace_b200 as pt
    from ipu_path_trace_b200.nif import write_nif_bin
    exe = os.path.join(ROOT, "host", "b200_path_trace")
    if not os.path.exists(exe):
        subprocess.run(["make", "-C", os.path.join(ROOT, "host"), "b200_path_trace"], check=True, capture_output=True)
    with tempfile.TemporaryDirectory() as tmp:
        assets = os.path.join(tmp, "assets.extra")
        os.makedirs(assets)
        meta = os.path.join(ROOT, "tests", "golden", "nif_metadata.txt")        # the shipped model's metadata, verbatim
        open(os.path.join(assets, "nif_metadata.txt"), "w").write(open(meta).read())
        write_nif_bin(os.path.join(assets, "nif.bin"), pt.synthetic_model())
        out = os.path.join(tmp, f"{tag}.png")
        cmd = [exe, "-o", out, "--assets", assets, "-w", str(width), "-h", str(height), "-s", str(spp), "--samples-per-step", str(sps),
               "--ipus", str(n_gpus), "--save-interval", str(save_interval)]
        # torch.distributed.run exports OMP_NUM_THREADS=1 to its ranks; the C++ host is its own program with a parallel
        # PNG / EXR encoder and must see the machine as a user's shell would give it.
        env = {k: v for k, v in os.environ.items() if k not in ("OMP_NUM_THREADS", "MKL_NUM_THREADS")}
        t0 = time.perf_counter()
        r = subprocess.run(cmd, capture_output=True, text=True, timeout=timeout, env=env)
        wall = time.perf_counter() - t0
        log = r.stderr
        if os.environ.get("B200PT_BENCH_LOGDIR"):
            open(os.path.join(os.environ["B200PT_BENCH_LOGDIR"], f"cli_{tag}_n{n_gpus}.log"), "w").write(log)
        if r.returncode != 0:
            return {"error": log[-400:], "cmd": " ".join(cmd[:1] + cmd[5:])}
        steps = [(float(a), float(b)) for a, b in re.findall(r"Completed render step \d+/\d+ in ([0-9.e+-]+) seconds \(Samples/sec ([0-9.e+-]+)\)", log)]
        final = re.findall(r"\] Samples/sec: ([0-9.e+-]+)", log)
        init = re.findall(r"Initialisation took ([0-9.e+-]+) seconds", log)
        acq = re.findall(r"Device acquisition took ([0-9.e+-]+) seconds", log)
        secs = np.array([a for a, _ in steps])
        rates = np.array([b for _, b in steps])
        files_ok = os.path.exists(out) and os.path.exists(out[:-4] + ".exr")
        return {"cmd": "b200_path_trace " + " ".join(cmd[5:]), "steps": len(steps), "run_level_samples_per_s": float(final[-1]) if final else None,
                "steady_state_samples_per_s": float(np.median(rates)) if len(rates) else None,
                "run_level_over_steady_state": (float(final[-1]) / float(np.median(rates))) if final and len(rates) else None,
                "initialisation_s": float(init[-1]) if init else None, "device_acquisition_s": float(acq[-1]) if acq else None, "process_wall_s": wall,
                "clock": "run-level Samples/sec as the reference computes it: from the start of execute() (after device attach / engine load, here after context + NCCL creation) to the end of the render",
                "slowest_step_over_median": float(secs.max() / np.median(secs)) if len(secs) else None,
                "saves": log.count("Saved images at step"), "png_and_exr_written": files_ok}


def multi_gpu_check(pt, model, dist, rank: int, world: int, local: int):
    """The reference's replicas render ONE image (src/PathTracerApp.cpp:205-252): a small image rendered sharded over the
    `world` ranks and reduced over NCCL must be the image one GPU renders whole. tests/test_gpu_multi.py asserts this, but
    needs a multi-GPU lease; this puts the same comparison where an N-GPU run actually happens. Collective: every rank calls."""
    import torch
    Wc, Hc, sps, steps = 208, 160, 8, 3

    def render(worklist):
        t = pt.PathTracer(Wc, Hc, device=local, max_worklist=len(worklist), nif_impl=pt.NIF_TCGEN05)
        t.init_nif_weights(model)
        t.init_render_settings(seed=1, samples_per_step=sps)
        t.setup(worklist)
        tot = [0, 0, 0]
        for _ in range(steps):
            t.path_trace(sync=False)
            t.film_accumulate()
            c = t.counters()
            tot[0] += c.paths; tot[1] += c.escaped; tot[2] += c.path_segments
        film = torch.zeros(Hc * Wc * 3, dtype=torch.float32, device="cuda")
        t.film_read_device(film.data_ptr(), 1.0 / steps)
        t.close()
        return film, tot

    shard = pt.shard_tiles(Wc, Hc, world)[rank]
    film, tot = render(shard)
    dist.reduce(film, dst=0)                                   # disjoint pixel support: the sum is a gather
    counts = torch.tensor(tot + [len(shard)], dtype=torch.int64, device="cuda")
    per_rank = [torch.zeros_like(counts) for _ in range(world)]
    dist.all_gather(per_rank, counts)
    torch.cuda.synchronize()
    if rank != 0:
        return None
    want, want_tot = render(pt.create_work_list_for_image(Wc, Hc))
    got_tot = [int(sum(int(x[i]) for x in per_rank)) for i in range(3)]
    a, b = film.double(), want.double()
    rel = ((a - b).abs() / b.abs().clamp_min(1e-7)).max().item()
    return {"image": f"{Wc}x{Hc}, {sps * steps} spp, sharded over {world} ranks + NCCL reduce vs whole on rank 0",
            "max_rel_diff": rel, "bit_equal": bool(torch.equal(film, want)),
            "counters_equal": got_tot == [int(x) for x in want_tot], "paths": got_tot[0], "escaped": got_tot[1], "path_segments": got_tot[2],
            "items_per_rank": [int(x[3]) for x in per_rank]}


# ---------------------------------------------------------------------------------------------
# GPU arm
# ---------------------------------------------------------------------------------------------
def run_ours(args) -> None:
    import torch
    import ipu_path_trace_b200 as pt

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the hot path has no CPU fallback (use --impl reference for the CPU arm)")
    torch.cuda.set_device(local)
    dist = None
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    # host-side barrier for the extras: an NCCL barrier would leave a spinning kernel on every idle rank's GPU, time-slicing
    # against the C++ host process that renders on those GPUs meanwhile
    cpu_group = dist.new_group(backend="gloo") if dist is not None else None

    def host_barrier():
        torch.cuda.synchronize()
        if dist is not None:
            dist.barrier(group=cpu_group)

    def barrier():
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()

    sps = args.samples_per_step
    model = pt.synthetic_model()
    worklist = pt.shard_tiles(W, H, world)[rank]
    if os.environ.get("B200PT_BENCH_SHARDS"):      # diagnosis only: one GPU renders the share it would have among N
        worklist = pt.shard_tiles(W, H, int(os.environ["B200PT_BENCH_SHARDS"]))[0]
    tracer = pt.PathTracer(W, H, device=local, max_worklist=len(worklist), nif_impl=pt.NIF_TCGEN05)
    tracer.init_nif_weights(model)
    tracer.init_render_settings(seed=1, samples_per_step=sps)
    tracer.setup(worklist)
    flops_per_eval = tracer.nif_flops_per_eval()
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")     # > 126 MB L2
    film_dev = torch.zeros(H * W * 3, dtype=torch.float32, device="cuda")

    def step_resident():
        tracer.path_trace(sync=False)
        tracer.film_accumulate()
        tracer.sync()

    # e2e: the reference-facing programs with HOST buffers every step, laid out like the reference's render loop
    # (PathTracerApp.cpp:690-709): page-locked work list and result buffers (b200pt_host_alloc), and the NEXT step's work
    # list staged with setup_next while this step's path_trace runs (the reference's double-buffered work list).
    pin_in = pt.PinnedRecords(len(worklist))
    pin_out = pt.PinnedRecords(len(worklist))
    pin_in.array[:] = worklist
    out_records = pin_out.array

    def step_e2e():
        tracer.path_trace(sync=False)             # "path_trace" (uses the list staged by the previous step's setup_next)
        tracer.setup_next(pin_in.array)           # H->D TraceRecords of the next step ("setup"), overlapped with the trace
        return tracer.read_results(out_records)   # D->H TraceRecords + counters ("read_results")

    def gather_film():
        """One reduce of the per-GPU films (disjoint pixel support, so the sum is a gather) to rank 0."""
        tracer.film_read_device(film_dev.data_ptr(), 1.0)
        if dist is not None:
            dist.reduce(film_dev, dst=0)
        torch.cuda.synchronize()

    for _ in range(args.warmup):
        step_resident()
        flush.fill_(1)
    barrier()

    clocks = ClockSampler(local)
    if rank == 0:
        clocks.start()
    # ---- timed: K resident steps and K end-to-end steps, in alternating blocks (resident, e2e, resident, e2e) so that
    # both see the same thermal / power state of a board that sits at its power cap (measured: whichever loop runs second
    # is ~1-2 % slower at 36 ms steps). Every step is bracketed by barrier + synchronize on both sides, L2 flushed before it.
    total = e2e_total = 0.0
    nif_ms = trace_ms = dev_ms = 0.0
    escaped = paths = launches = 0
    blocks = 2 if args.steps >= 4 else 1
    sizes = [args.steps // blocks + (1 if b < args.steps % blocks else 0) for b in range(blocks)]
    for nb in sizes:
        tracer.setup(worklist)                    # resident block: work list in HBM, accumulators cleared by film_accumulate
        step_resident()
        for _ in range(nb):
            flush.fill_(1)                        # L2 flush between timed iterations
            barrier()
            t0 = time.perf_counter()
            step_resident()
            barrier()
            total += time.perf_counter() - t0
            c = tracer.counters()
            nif_ms += c.nif_ms; trace_ms += c.trace_ms; dev_ms += c.total_ms
            escaped += c.escaped; paths += c.paths; launches += c.kernel_launches + 1
        tracer.setup(pin_in.array)                # end-to-end block: host buffers every step
        step_e2e()                                # fills the pipeline (stages the next work list)
        for _ in range(nb):
            flush.fill_(1)
            barrier()
            t0 = time.perf_counter()
            step_e2e()
            barrier()
            e2e_total += time.perf_counter() - t0
    # ---- the per-save-interval gather (outside the sample loop) ----
    gather_film()
    barrier()
    t0 = time.perf_counter()
    gather_film()
    barrier()
    gather_ms = 1e3 * (time.perf_counter() - t0)
    clk = clocks.stop() if rank == 0 else None

    stats = torch.tensor([total, e2e_total, nif_ms, trace_ms, dev_ms, gather_ms], dtype=torch.float64, device="cuda")
    sums = torch.tensor([escaped, paths, launches], dtype=torch.float64, device="cuda")
    if dist is not None:
        dist.all_reduce(stats, op=dist.ReduceOp.MAX)
        dist.all_reduce(sums, op=dist.ReduceOp.SUM)
    total, e2e_total, nif_ms, trace_ms, dev_ms, gather_ms = stats.tolist()
    escaped, paths, launches = sums.tolist()

    extras = {}
    if rank == 0 and not args.no_extras:
        try:
            extras["c4"] = run_c4(tracer, flops_per_eval)
        except Exception as e:      # an extra must never lose the headline
            extras["c4"] = {"error": repr(e)[:300]}
    if dist is not None and not args.no_extras:
        try:
            chk = multi_gpu_check(pt, model, dist, rank, world, local)
        except Exception as e:
            chk = {"error": repr(e)[:300]}
        if rank == 0:
            extras["multi_gpu_check"] = chk
    tracer.close()
    tracer = None
    pin_in.close(); pin_out.close()
    del flush, film_dev
    torch.cuda.empty_cache()
    host_barrier()
    if rank == 0 and not args.no_extras:
        # the reference's README run, to completion, through the C++ host on the same GPUs (the ranks idle meanwhile)
        try:
            extras["cli_c2"] = run_cli(world, W, H, args.cli_spp, SAMPLES_PER_STEP, SAVE_INTERVAL, "c2", timeout=900)
            if world == 8:
                extras["cli_c5"] = run_cli(world, 3840, 2160, 6400, 64, 10, "c5", timeout=600)
                extras["cli_c5"]["workload"] = "C5: 3840x2160 progressive render, PNG+EXR every 10 steps of 64 samples, 8 GPUs"
        except Exception as e:
            extras["cli_c2"] = {"error": repr(e)[:300]}
    host_barrier()

    if rank == 0:
        samples_per_job_step = W * H * sps
        value = samples_per_job_step * args.steps / total
        e2e_value = samples_per_job_step * args.steps / e2e_total
        peak_tf, _, peak_src = measured_peaks()
        # roofline of the dominant kernel (NIF tcgen05): algorithmic flops of the launches of one rank
        # / the CUDA-event time of those launches on that rank's stream (max over ranks).
        flops_rank = flops_per_eval * escaped / world
        achieved_tf = flops_rank / (nif_ms * 1e-3) / 1e12 if nif_ms > 0 else 0.0
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": 1e3 * total / args.steps, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
            "dtype": "f16 x f16 -> f32 (tcgen05 MLP); f32 path tracing", "data":
            "synthetic (built-in sphere scene; seeded synthetic NIF weights in the shipped topology: converted.hdf5 is missing from the reference snapshot)",
            "config": {"workload": workload_name(sps), "sharding": f"{world} GPU(s), pixel tiles 16x16 shuffled and dealt round-robin",
                       "samples_per_job_step": samples_per_job_step, "l2": "flushed (256 MiB write) between timed steps",
                       "escaped_fraction": escaped / max(1.0, paths), "nif_flops_per_eval": flops_per_eval,
                       "device_ms_per_step": dev_ms / args.steps, "trace_ms_per_step": trace_ms / args.steps,
                       "nif_ms_per_step": nif_ms / args.steps, "film_gather_ms_per_save_interval": gather_ms,
                       "save_interval_steps": SAVE_INTERVAL},
            "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": int(20 * W * H), "d2h_bytes_per_step": int(20 * W * H),
                    "ms_per_step": 1e3 * e2e_total / args.steps},
            "gpu_launches": int(launches),
            "clocks": clk,
            "roofline": {"bound": "tensor", "achieved": achieved_tf, "peak": peak_tf, "unit": "TFLOP/s", "frac": achieved_tf / peak_tf,
                         "traffic": nif_traffic_bytes()[0], "traffic_source": nif_traffic_bytes()[1], "kernel": "nif_tc_kernel", "peak_source": peak_src,
                         "whole_step_frac": (flops_per_eval * escaped / world) / (dev_ms * 1e-3) / 1e12 / peak_tf if dev_ms > 0 else 0.0},
        }
        line["config"].update(extras)
        if world == 1 and not args.no_cpu_baseline:
            v, desc, cores = cpu_sample_run(15.0)
            line["cpu_baseline"] = {"value": v, "unit": UNIT, "cores": cores, "kind": "port", "sample": desc}
        print(json.dumps(line), flush=True)
    if dist is not None:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--samples-per-step", type=int, default=SAMPLES_PER_STEP)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--cpu-seconds", type=float, default=0.0, help="--impl reference: seconds of CPU work per step (default: sized so the run takes ~2 minutes)")
    ap.add_argument("--no-extras", action="store_true", help="skip the C4 microbenchmark and the C++ host runs (config.c4 / cli_c2 / cli_c5)")
    ap.add_argument("--cli-spp", type=int, default=100000, help="samples per pixel of the C++ host run (README.md:82 uses 100000)")
    args = ap.parse_args()
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if args.impl == "reference":
        run_reference(args)
        return
    if args.gpus != world:
        if world == 1 and args.gpus > 1:
            # convenience: re-launch under torchrun (the driver launches torchrun itself)
            cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={args.gpus}",
                   "--master-addr", "127.0.0.1", "--master-port", "29517", os.path.abspath(__file__)] + sys.argv[1:]
            raise SystemExit(subprocess.call(cmd))
        raise SystemExit(f"--gpus {args.gpus} does not match WORLD_SIZE {world}")
    run_ours(args)


if __name__ == "__main__":
    main()
