#!/usr/bin/env python
"""bench.py -- FProxGD iterations/s on the BASELINE.json workload, with roofline, parity and CPU baseline.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]
    python -m torch.distributed.run --nproc-per-node N ... bench.py --gpus N ...

Our arm.  A "step" is one drop-in solve of the hot path (gl_FProxGD_primal semantics) with `maxit`
iterations per continuation stage (3*maxit iterations, no early stop possible for maxit <= 70) on the
synthetic workload  m=32768 n=131072 l=2 fp32 (BASELINE.json configs[3]; ~17 GB of A, column-group
sharded over the ranks).  value = iterations/s of the whole job with A resident in HBM; e2e = the same
metric through the host-buffer entry point (NumPy in, NumPy out: upload of A, b, x0 and download of x
inside the timed region).  Beside the headline (outside its timed region): `parity` (the same solve with
the CPU sample's iteration budget, objective compared with the CPU run), `extra` (BASELINE configs[2] and
[4] and ProxGD on the headline problem, each with its roofline), `time_to_tol`, `homework_fixture`.

Reference arm (--impl reference).  The reference's own CPU implementation of the path -- the unmodified
solver files staged under oracle/_ref by oracle/make_ref.py, else the oracle port -- on the box's host
cores, same metric / config, every step a bounded sample of the workload.  It never loads the product
library: the problem comes from the CPU restatement of the generator (oracle/gen.py), which is
bit-identical to the device generator.  Rank 0 alone works; torchrun's OMP_NUM_THREADS=1 is undone first.

One JSON line is printed by rank 0.
"""
import os
import sys


def _host_cores():
    try:
        return len(os.sched_getaffinity(0))
    except AttributeError:
        return os.cpu_count() or 1


def _reference_arm(argv):
    for i, a in enumerate(argv):
        if a == "--impl" and i + 1 < len(argv) and argv[i + 1] == "reference":
            return True
        if a == "--impl=reference":
            return True
    return False


if _reference_arm(sys.argv):
    # torch.distributed.run exports OMP_NUM_THREADS=1 to every rank; the CPU arm wants every host core.
    # Must happen before NumPy (OpenBLAS) and libgomp are loaded.
    for _v in ("OMP_NUM_THREADS", "OPENBLAS_NUM_THREADS", "MKL_NUM_THREADS"):
        os.environ[_v] = str(_host_cores())

import argparse  # noqa: E402
import ctypes as C  # noqa: E402
import json  # noqa: E402
import math  # noqa: E402
import subprocess  # noqa: E402
import threading  # noqa: E402
import time  # noqa: E402

import numpy as np  # noqa: E402

ROOT = os.path.dirname(os.path.abspath(__file__))
for p in (ROOT, os.path.join(ROOT, "convex-optimization_b200")):
    if p not in sys.path:
        sys.path.insert(0, p)

METRIC = "FProxGD iterations/s"
UNIT = "it/s"
CPU_SAMPLE_MAXIT = 2          # iterations per stage of every CPU sample and of the GPU parity solve
REFERENCE_BUDGET_S = 100.0    # wall-clock target of the reference arm's K + W steps


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    # --rows / --cols: spellings to use under torchrun, whose own parser claims the prefix "--m"
    ap.add_argument("--m", "--rows", dest="m", type=int, default=32768)
    ap.add_argument("--n", "--cols", dest="n", type=int, default=131072)
    ap.add_argument("--l", "--rhs", dest="l", type=int, default=2)
    ap.add_argument("--dtype", default="f32", choices=["f32", "f64"])
    ap.add_argument("--algo", default="FProxGD", choices=["FProxGD", "ProxGD"])
    ap.add_argument("--maxit", type=int, default=50, help="iterations per continuation stage in one step")
    ap.add_argument("--seed", type=int, default=97006855)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-extra", action="store_true", help="skip the C3 / C5 / ProxGD side measurements")
    ap.add_argument("--exchange", default="peer", choices=["peer", "nccl"],
                    help="multi-GPU exchange: peer-mailbox push fused into the kernels (product) or the ncclAllReduce baseline")
    ap.add_argument("--tensor-cores", default="auto", choices=["auto", "3xtf32", "tf32+bf16", "f16x2", "tf32", "off"],
                    help="wide multi-RHS mode (l >= 16): product scheme on the tensor cores")
    ap.add_argument("--no-converge", action="store_true", help="skip the run-to-convergence / time-to-1e-8 report")
    return ap.parse_args()


def mp_step(m, n):
    """Marchenko-Pastur bound on 1/lambda_max(A^T A); the reference defaults diverge off the fixture."""
    return float(1.0 / (math.sqrt(m) + math.sqrt(n)) ** 2)


def bench_config(args):
    """`config` of the JSON line: a function of the command line only, so both arms print the same object."""
    es = 4 if args.dtype == "f32" else 8
    return {"workload": "group-lasso %s m=%d n=%d l=%d %s, mu=1e-2, alpha0=1/(sqrt(m)+sqrt(n))^2; A ~ N(0,1) from the "
                        "counter-based stream of seed %d, u with 10%% non-zero row groups, b = A u, x0 ~ N(0,1)" % (
                            args.algo, args.m, args.n, args.l, args.dtype, args.seed),
            "l2_policy": "A (%.1f GB over %d GPU(s)) is far larger than the 126 MB L2; no flush needed" % (
                args.m * args.n * es / 1e9, args.gpus)}


def mem_available_gb():
    try:
        for line in open("/proc/meminfo"):
            if line.startswith("MemAvailable:"):
                return int(line.split()[1]) / 1e6
    except OSError:
        pass
    return 0.0


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled during the timed region (B200_PROFILING.md)."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,"
         "clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index):
        self.idx = gpu_index
        self.proc = None
        self.lines = []

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.idx), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits", "-lms", "200"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thr = threading.Thread(target=self._pump, daemon=True)
            self.thr.start()
        except OSError:
            self.proc = None

    def _pump(self):
        for ln in self.proc.stdout:
            self.lines.append((time.time(), ln.strip()))

    def stop(self, t0=None, t1=None):
        """Summary of the samples taken in [t0, t1] (the timed region); if the region was shorter than
        the sampling period, of all samples since start() (warm-up runs the same kernels)."""
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except subprocess.TimeoutExpired:
            self.proc.kill()
        sm, smax, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        inside = [ln for (ts, ln) in self.lines if t0 is not None and t0 <= ts <= (t1 or ts) + 0.25]
        window = "timed region" if inside else "warm-up + timed region"
        for ln in (inside or [ln for (_, ln) in self.lines]):
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 9:
                continue
            try:
                sm.append(float(f[1])); smax.append(float(f[2]))
            except ValueError:
                continue
            for nm, val in zip(names, f[5:9]):
                if val.lower().startswith("active"):
                    reasons.add(nm)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(smax) if smax else None,
                "samples": len(sm), "window": window, "reasons": sorted(reasons)}


def dist_setup(args):
    import torch
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world > 1:
        import torch.distributed as dist
        torch.cuda.set_device(local)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    else:
        torch.cuda.set_device(0)
    return rank, local, world


def make_problem(m, n, l, seed, dtype, rank, world):
    """Synthetic problem of code/main.py:37-51 at scale (glprox.dataset.gen_data_device): A ~ N(0,1) from
    the counter-based device generator (identical values for any sharding), 10% non-zero row groups in
    u, b = A u."""
    from glprox import dataset
    A, b, x0_loc, _u, mu, _ = dataset.gen_data_device(m, n, l, seed, dtype, rank, world)
    return A, b, x0_loc, mu


def load_peaks():
    try:
        return json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except (OSError, ValueError):
        return {}


def device_max(ms, world):
    if world == 1:
        return ms
    import torch
    import torch.distributed as dist
    t = torch.tensor([ms], device="cuda", dtype=torch.float64)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return t.item()


def barrier(world):
    import torch
    if world > 1:
        import torch.distributed as dist
        dist.barrier()
    torch.cuda.synchronize()


def timed_solves(prob, algo, x0, mu, opts, steps, warmup, world, sampler=None):
    """W untimed solves, then K solves between barrier + synchronize, CUDA events on the solve stream, max
    over ranks.  The library's own per-launch CUDA-event profile is switched on for the timed solves."""
    import torch
    from glprox import _lib
    lib = _lib.load()
    for _ in range(warmup):
        x, k, out = prob.solve(algo, x0, mu, opts)
    _lib.check(lib.glp_set_profiling(prob.handle, 1))
    barrier(world)
    t_wall0 = time.time()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ev0.record()
    iters = launches = 0
    for _ in range(steps):
        x, k, out = prob.solve(algo, x0, mu, opts)
        iters += k
        launches += out["glprox"]["n_launches"]
    ev1.record()
    barrier(world)
    t_wall1 = time.time()
    ms = device_max(ev0.elapsed_time(ev1), world)
    prof = _lib.GlpProfile()
    _lib.check(lib.glp_get_profile(prob.handle, C.byref(prof)))
    _lib.check(lib.glp_set_profiling(prob.handle, 0))
    return {"iters": iters, "ms": ms, "launches": launches, "out": out, "prof": prof, "wall": (t_wall0, t_wall1)}


def ncu_traffic(dtype, m, n, l, world):
    """dram__bytes_read.sum + dram__bytes_write.sum per launch of the dominant kernel, from the committed
    `ncu --set full` capture of this workload (profiles/ncu_traffic.json names the .csv); None if not captured."""
    try:
        cap = json.load(open(os.path.join(ROOT, "profiles", "ncu_traffic.json")))
    except (OSError, ValueError):
        return None
    e = cap.get("%s_m%d_n%d_l%d_g%d" % (dtype, m, n, l, world))
    return e if isinstance(e, dict) else None


def stream_roofline(prof, value, peaks, cap):
    """l = 2: the two streams of A (K1 forward, K2 transposed) against the measured HBM copy peak."""
    peak = float(peaks.get("hbm_gbs", 6650.0))
    n_str = prof.fwd_launches + prof.tpass_launches
    avg_ms = (prof.fwd_ms_sum + prof.tpass_ms_sum) / max(n_str, 1)
    achieved = prof.bytes_per_launch / (avg_ms * 1e-3) / 1e9 if n_str else 0.0
    per = lambda cnt, ms: {"launches": int(cnt), "avg_ms": ms / max(cnt, 1),
                           "GBps": prof.bytes_per_launch * cnt / max(ms, 1e-9) / 1e6}
    return {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
            "traffic": (cap or {}).get("bytes_per_launch"),
            "traffic_source": (cap or {}).get("source", None),
            "peak_source": "MEASURED_PEAKS.json hbm_gbs (copy, read+write)" if peaks else "fallback 6650 GB/s",
            "kernels": {"glp_fwd_kernel": per(prof.fwd_launches, prof.fwd_ms_sum),
                        "glp_tpass_kernel": per(prof.tpass_launches, prof.tpass_ms_sum)},
            "bytes_per_launch": prof.bytes_per_launch,
            "whole_iteration_GBps": 2.0 * prof.bytes_per_launch * value / 1e9}


def wide_roofline(prof, value, peaks, cap, m, n_local, l, tensor_cores):
    """l >= 16 (fp32): both products run on tcgen05; report the binding roofline of the two (HBM stream of the
    resident operand set vs tensor pipe in TF32-product units) and the other one beside it."""
    peak = float(peaks.get("hbm_gbs", 6650.0))
    tc_mode = "f16x2" if tensor_cores == "auto" else tensor_cores   # auto = fp16 pairs (finite A)
    # tensor time in TF32-product units: 3xTF32 issues 3 TF32 products; tf32+bf16 one TF32 product and two bf16
    # products at twice the rate (= 2 units); fp16 pairs three kind::f16 products at twice the rate (= 1.5)
    passes = {"tf32": 1.0, "tf32+bf16": 2.0, "f16x2": 1.5}.get(tc_mode, 3.0)
    n_it = max(int(prof.tpass_launches), 1)
    # operand columns on the tensor cores per iteration: l (A.xn) + l (A^T r); the product of the thresholded
    # iterate comes from A.xn by a sparse CUDA-core correction (csrc/glp_thrfix.cuh) unless switched off
    thrfix = os.environ.get("GLPROX_WIDE_THRFIX", "1") != "0"
    op_cols = (2 if thrfix else 3) * l
    flops = passes * 2.0 * m * n_local * op_cols * n_it
    t_s = (prof.fwd_ms_sum + prof.tpass_ms_sum) * 1e-3
    tf = flops / max(t_s, 1e-12) / 1e12
    tpeak = float(peaks.get("bf16_tflops", 1590.0)) / 2.0
    stream_x = 1.0 if tc_mode in ("tf32", "f16x2") else 2.0   # bytes of the resident set per fp32 element / 4
    stream_bytes = stream_x * prof.bytes_per_launch * (prof.fwd_launches + prof.tpass_launches)
    hbm_stream = stream_bytes / max(t_s, 1e-12) / 1e9
    t_min_tensor, t_min_hbm = flops / (tpeak * 1e12), stream_bytes / (peak * 1e9)
    tensor_part = {"achieved": tf, "peak": tpeak, "unit": "TFLOP/s", "frac": tf / tpeak,
                   "peak_source": "MEASURED_PEAKS.json bf16_tflops / 2 (TF32-product units)"}
    hbm_part = {"achieved": hbm_stream, "peak": peak, "unit": "GB/s", "frac": hbm_stream / peak,
                "peak_source": "MEASURED_PEAKS.json hbm_gbs (copy, read+write)" if peaks else "fallback 6650 GB/s"}
    bound = "hbm" if t_min_hbm > t_min_tensor else "tensor"
    wide_traffic = (cap or {}).get("bytes_per_launch") if (cap or {}).get("tensor_cores") == tc_mode else None
    return {"bound": bound, **(hbm_part if bound == "hbm" else tensor_part), "traffic": wide_traffic,
            "traffic_source": (cap or {}).get("source") if wide_traffic else None,
            "other_roofline": {"bound": "tensor" if bound == "hbm" else "hbm",
                               **(tensor_part if bound == "hbm" else hbm_part)},
            "min_ms_per_iteration": {"tensor": 1e3 * t_min_tensor / n_it, "hbm": 1e3 * t_min_hbm / n_it,
                                     "measured_products": 1e3 * t_s / n_it},
            "whole_iteration_frac_of_hbm": stream_x * 2.0 * prof.bytes_per_launch * value / 1e9 / peak,
            "tensor_cores": tc_mode, "tensor_operand_columns_per_iteration": op_cols,
            "forward_product": ("A.xn on tcgen05 (l columns) + sparse CUDA-core correction for A.thr(xn), both inside the "
                                "fwd window" if thrfix else "A.[xn | thr(xn)] on tcgen05 (2l columns)"),
            "kernels": {"glp_gemm_tf32_kernel<fwd>": {"launches": int(prof.fwd_launches),
                                                      "avg_ms": prof.fwd_ms_sum / max(prof.fwd_launches, 1)},
                        "glp_gemm_tf32_kernel<tpass>": {"launches": int(prof.tpass_launches),
                                                        "avg_ms": prof.tpass_ms_sum / max(prof.tpass_launches, 1)}}}


def exchange_text(args, world):
    if world == 1:
        return "none (1 GPU)"
    if args.exchange == "nccl":
        return ("BASELINE: two ncclAllReduce calls per forward stream (MAX of one double, SUM of the m x 2l "
                "partial product + 8 scalars) on the solver stream")
    if args.l >= 16 and args.dtype == "f32":
        return ("reduce-scatter + all-gather of the m x 2l partial products over row slices, pushed through peer "
                "mailboxes over NVLink (3 small kernels per forward product); no NCCL on the data path")
    return ("m x 2l partial products + 5 scalars pushed into peer mailboxes over NVLink from inside the forward "
            "kernel; rank-ordered sum on every rank; no NCCL on the data path")


def run_ours(args):
    import torch
    from glprox import GroupLassoProblem, _lib
    rank, local, world = dist_setup(args)
    _lib.load()
    if world != args.gpus:
        raise SystemExit("--gpus %d but WORLD_SIZE=%d (launch with torch.distributed.run)" % (args.gpus, world))
    A, b, x0, mu = make_problem(args.m, args.n, args.l, args.seed, args.dtype, rank, world)
    opts = {"maxit": args.maxit, "alpha0": mp_step(args.m, args.n)}
    if args.tensor_cores != "auto":
        opts["tensor_cores"] = args.tensor_cores
    prob = GroupLassoProblem(A, b)
    if world > 1:
        from glprox import attach_distributed
        attach_distributed(prob, args.n, exchange=args.exchange)

    # ---- device-resident throughput (the headline)
    sampler = ClockSampler(local)
    sampler.start()
    r = timed_solves(prob, args.algo, x0, mu, opts, args.steps, args.warmup, world)
    t_wall0, t_wall1 = r["wall"]
    if t_wall1 - t_wall0 < 0.5:
        time.sleep(0.3)  # let one more nvidia-smi sample land
    clocks = sampler.stop(t_wall0, t_wall1)
    ms, iters, prof = r["ms"], r["iters"], r["prof"]
    value = iters / (ms / 1e3)
    peaks = load_peaks()
    cap = ncu_traffic(args.dtype, args.m, args.n, args.l, world)
    if args.l >= 16 and args.dtype == "f32":
        roofline = wide_roofline(prof, value, peaks, cap, args.m, A.shape[1], args.l, args.tensor_cores)
    else:
        roofline = stream_roofline(prof, value, peaks, cap)
    line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms / args.steps, "higher_is_better": True,
            "scaling": "strong", "vs_baseline": None, "dtype": args.dtype, "data": "synthetic",
            "config": bench_config(args),
            "parallel": {"sharding": "column groups of A / row groups of x over %d rank(s)" % world,
                         "exchange": exchange_text(args, world)},
            "maxit_per_stage": args.maxit, "iterations_per_step": iters // max(args.steps, 1),
            "final_objective": float(r["out"]["fval"]),
            "gpu_launches": int(r["launches"]), "clocks": clocks, "roofline": roofline}
    return line, (prob, A, b, x0, mu, opts, rank, world)


def run_parity_gpu(args, ctx):
    """The GPU half of `parity`: the same solver, same inputs, with the CPU sample's iteration budget."""
    prob, A, b, x0, mu, opts, rank, world = ctx
    o = dict(opts, maxit=CPU_SAMPLE_MAXIT)
    x, k, out = prob.solve(args.algo, x0, mu, o)
    return {"maxit": CPU_SAMPLE_MAXIT, "iters": int(k), "gpu_fval": float(out["fval"]),
            "gpu_f_hist": [float(v) for v in out["f_hist"]], "bar": 1e-5 if args.dtype == "f32" else 1e-10,
            "cpu_fval": None, "rel_diff": None}


def run_e2e(args, ctx, steps):
    """Same metric through the host-buffer entry point: NumPy arrays in pinned host memory go in,
    a NumPy x comes out; the upload of A, b, x0 and the download of x are inside the timed region."""
    import torch
    from gl_FProxGD_primal import gl_FProxGD_primal
    from gl_ProxGD_primal import gl_ProxGD_primal
    prob, A, b, x0, mu, opts, rank, world = ctx
    need_gb = A.numel() * A.element_size() / 1e9
    if mem_available_gb() < 1.5 * need_gb + 4:
        return None, "host RAM too small for a %.1f GB host copy of A" % need_gb
    fn = gl_FProxGD_primal if args.algo == "FProxGD" else gl_ProxGD_primal
    A_h = torch.empty(A.shape, dtype=A.dtype, pin_memory=True)
    A_h.copy_(A)
    b_h = b.cpu().pin_memory()
    x0_h = x0.cpu().pin_memory()
    torch.cuda.synchronize()
    An, bn, x0n = A_h.numpy(), b_h.numpy(), x0_h.numpy()
    opts = dict(opts)
    if world > 1:
        import torch.distributed as dist
        opts["dist"] = True  # every rank passes its own column block of A / rows of x0
    fn(x0n, An, bn, mu, dict(opts))  # warm-up (page-in, bounce buffers)
    iters = 0
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(steps):
        x, k, out = fn(x0n, An, bn, mu, dict(opts))
        iters += k
    torch.cuda.synchronize()
    dt = time.perf_counter() - t0
    h2d = An.nbytes + bn.nbytes + x0n.nbytes
    d2h = x.nbytes + 8 * k
    h2d_rank = h2d
    if world > 1:
        t = torch.tensor([dt, float(h2d), float(d2h)], device="cuda", dtype=torch.float64)
        dist.all_reduce(t[:1], op=dist.ReduceOp.MAX)
        dist.all_reduce(t[1:])
        dt, h2d, d2h = t[0].item(), t[1].item(), t[2].item()
    return {"value": iters / dt, "unit": UNIT, "h2d_bytes_per_step": int(h2d), "d2h_bytes_per_step": int(d2h),
            "ms_per_step": 1e3 * dt / steps, "final_objective": float(out["fval"]),
            "h2d_bytes_per_rank_per_step": int(h2d_rank),
            "upload_ms_per_step": out["glprox"].get("upload_ms"),
            "call_breakdown_ms": {k: out["glprox"].get(k) for k in ("upload_ms", "attach_ms", "solve_ms", "teardown_ms")
                                  if out["glprox"].get(k) is not None} or None,
            "api": "gl_%s_primal(x0, A, b, mu, opts) with NumPy arrays (pinned host memory)" % args.algo}, (An, bn, x0n)


def time_to_tol(f_hist, f_ref, seconds, tol=1e-8):
    """First k with |f_hist[k] - f_ref| / |f_ref| <= tol, and the wall-clock share of the solve up to
    it (iterations cost the same; SURVEY.md 8(d) 'time-to-1e-8')."""
    fh = np.asarray(f_hist, dtype=np.float64)
    hit = np.flatnonzero(np.abs(fh - f_ref) <= tol * abs(f_ref))
    if hit.size == 0 or fh.size == 0:
        return None, None
    k = int(hit[0]) + 1
    return k, seconds * k / fh.size


def run_convergence(args, ctx):
    """Second half of BASELINE.json's metric ("time-to-1e-8 objective") on the headline workload: ONE
    solve with the reference's default iteration budget (maxit per stage from default_opts, all three
    continuation stages, the reference's stable-length stop rule), device-resident; reports the
    iteration and the time at which the objective first comes within 1e-8 (relative) of its final
    value.  Outside the timed region of the headline number."""
    prob, A, b, x0, mu, opts, rank, world = ctx
    from glprox import DEFAULT_OPTS
    full = dict(opts)
    full["maxit"] = DEFAULT_OPTS[args.algo]["maxit"]
    x, k, out = prob.solve(args.algo, x0, mu, full)
    secs = device_max(out["glprox"]["device_ms"], world) * 1e-3
    fh = np.asarray(out["f_hist"], dtype=np.float64)
    k8, s8 = time_to_tol(fh, float(fh[-1]), secs, 1e-8)
    k6, s6 = time_to_tol(fh, float(fh[-1]), secs, 1e-6)
    return {"opts": "reference defaults (maxit=%d per stage, ftol=1e-6, stable_len_threshold=70) with alpha0=1/(sqrt(m)+sqrt(n))^2" % full["maxit"],
            "iters": int(k), "seconds": secs, "it_per_s": k / secs, "final_objective": float(out["fval"]),
            "stage_end": out["glprox"]["stage_end"], "sparsity": out["glprox"]["sparsity"],
            "k_to_1e-6_of_final": k6, "s_to_1e-6_of_final": s6,
            "k_to_1e-8_of_final": k8, "s_to_1e-8_of_final": s8,
            "reached_before_budget": bool(k8 is not None and k8 < k),
            "note": "when the iteration budget ends before the stop rule fires (k_to == iters) the objective is still "
                    "decreasing: fp32 trajectories of such un-converged runs decorrelate at the 1e-4 level between "
                    "summation orders (1/2/4/8 shards), as they do between NumPy fp32 and fp64 (SURVEY.md 8(c))"}


def run_extra(args, ctx):
    """Side measurements the driver would otherwise never see (outside the headline's timed region), at every N:
    BASELINE configs[2] (C3, 8192 x 32768, l = 2), configs[4] (C5, 16384 x 65536, l = 128 on tcgen05, default
    operand set) and ProxGD on the headline problem.  Each: W = 1 solve, K = 2 timed solves of maxit = 10."""
    import torch
    from glprox import GroupLassoProblem, attach_distributed
    prob, A, b, x0, mu, opts, rank, world = ctx
    peaks = load_peaks()
    res = {}

    def one(name, pr, algo, x0_, mu_, o, m, n, n_loc, l, wide, note):
        try:
            r = timed_solves(pr, algo, x0_, mu_, o, 2, 1, world)
            v = r["iters"] / (r["ms"] / 1e3)
            cap = ncu_traffic("f32", m, n, l, world)
            rf = wide_roofline(r["prof"], v, peaks, cap, m, n_loc, l, "auto") if wide else stream_roofline(r["prof"], v, peaks, cap)
            res[name] = {"workload": note, "value": v, "unit": UNIT, "iters_timed": r["iters"],
                         "ms_per_iteration": r["ms"] / max(r["iters"], 1), "final_objective": float(r["out"]["fval"]),
                         "gpu_launches": int(r["launches"]), "n_fwd": int(r["out"]["glprox"]["n_fwd"]),
                         "n_tpass": int(r["out"]["glprox"]["n_tpass"]), "roofline": rf}
        except Exception as e:  # noqa: BLE001  (secondary report; never lose the headline line)
            res[name] = {"workload": note, "error": repr(e)}

    if args.dtype == "f32" and args.l == 2:
        one("proxgd_c4", prob, "ProxGD", x0, mu, dict(opts), args.m, args.n, A.shape[1], args.l, False,
            "gl_ProxGD_primal on the headline problem, maxit=%d per stage" % args.maxit)
    for name, (m, n, l, wide) in (("c3", (8192, 32768, 2, False)), ("c5", (16384, 65536, 128, True))):
        note = "group-lasso FProxGD m=%d n=%d l=%d f32, maxit=%d per stage%s" % (
            m, n, l, args.maxit, ", wide multi-RHS mode on tcgen05 (fp16 pairs, CTA-pair kernel)" if wide else "")
        try:
            A2, b2, x02, mu2 = make_problem(m, n, l, args.seed, "f32", rank, world)
            p2 = GroupLassoProblem(A2, b2)
            if world > 1:
                attach_distributed(p2, n, exchange=args.exchange)
        except Exception as e:  # noqa: BLE001
            res[name] = {"workload": note, "error": repr(e)}
            continue
        one(name, p2, "FProxGD", x02, mu2, {"maxit": args.maxit, "alpha0": mp_step(m, n)}, m, n, A2.shape[1], l, wide, note)
        p2.close()
        del A2, b2, x02, p2
        torch.cuda.empty_cache()
    # gl_Admm_dual at a streaming size: C3's matrix, blocked set-up + two streams of A and two
    # of Y = L^-T per iteration; rank 0 only (the dual solver is single-GPU)
    if rank == 0 and args.dtype == "f32":
        note = "gl_Admm_dual m=8192 n=32768 l=2 f32, maxit=20 (Gram on tcgen05, blocked fp64 Cholesky, z = Y(Y^T w) as streams)"
        try:
            m, n, l = 8192, 32768, 2
            A2, b2, x02, mu2 = make_problem(m, n, l, args.seed, "f32", 0, 1)
            p2 = GroupLassoProblem(A2, b2)
            best = None
            for _ in range(2):
                xa, ka, oa = p2.solve_admm_dual(x02, mu2, {"maxit": 20})
                best = oa["glprox"]["device_ms"] if best is None else min(best, oa["glprox"]["device_ms"])
            res["admm_c3"] = {"workload": note, "iters": int(ka), "ms_total": best, "final_objective": float(oa["fval"]),
                              "gpu_launches": int(oa["glprox"]["n_launches"]),
                              "note": "ms_total includes the factorisation (as the reference's call does, gl_ADMM_dual.py:57)"}
            p2.close()
            del A2, b2, x02, p2
            torch.cuda.empty_cache()
        except Exception as e:  # noqa: BLE001
            res["admm_c3"] = {"workload": note, "error": repr(e)}
    return res


def run_homework(args):
    """BASELINE.json configs[0]/[1]: the reference's own fixture (m=256 n=512 l=2 mu=1e-2, fp64, default
    opts, seed 97006855) on the persistent cluster kernel, next to the CPU oracle port timed in the same
    process.  Reported beside the headline line: it/s, seconds to reach 1e-8 of (i) the solver's own
    final objective (drop-in defaults) and (ii) the true optimum f* with opts={'thres': 1e-8}."""
    import torch
    from glprox import GroupLassoProblem
    from glprox.dataset import gen_data
    from oracle import glprox_oracle as orc   # CPU leg only: the port timed beside the GPU solves
    F_STAR = 6.1037724766016532e-01   # high-accuracy optimum; Mosek/Gurobi publish 6.10377E-01
    _, _, _, mu, A, b, u, x0 = gen_data(97006855, 512, 256, 2)
    prob = GroupLassoProblem(torch.from_numpy(A).cuda(), torch.from_numpy(b).cuda())
    x0d = torch.from_numpy(x0).cuda()
    res = {"workload": "Test_group_lasso fixture m=256 n=512 l=2 mu=1e-2 fp64, default opts",
           "published_optimum_mosek_gurobi": 6.10377e-01, "f_star": F_STAR}
    smem_bytes = 2.0 * A.nbytes   # K4 reads the resident A twice per iteration from shared memory
    for algo in ("FProxGD", "ProxGD"):
        row = {}
        for be in ("persistent", "stream"):
            best = None
            for _ in range(3):
                x, k, out = prob.solve(algo, x0d, mu, {"backend": be})
                ms = out["glprox"]["device_ms"]
                best = ms if best is None else min(best, ms)
            kk, tsec = time_to_tol(out["f_hist"], float(out["fval"]), best / 1e3)
            row[be] = {"iters": k, "fval": float(out["fval"]), "ms": best, "it_per_s": k / (best / 1e3),
                       "us_per_iteration": 1e3 * best / max(k, 1),
                       "k_to_1e-8_of_final": kk, "s_to_1e-8_of_final": tsec,
                       "launches": out["glprox"]["n_launches"], "n_fwd": int(out["glprox"]["n_fwd"])}
        # shared-memory roofline of the persistent kernel: each CTA of the cluster reads its slice of A twice per
        # iteration; B300_MICROARCH.md: 128 B/clk/SM of shared-memory bandwidth
        ctas = int(os.environ.get("GLPROX_CLUSTER", "0")) or 8
        smem_peak_gbs = ctas * 128.0 * 1.965   # bytes/clk/SM x GHz at the maximum SM clock
        row["persistent"]["smem_roofline_frac"] = (smem_bytes * row["persistent"]["it_per_s"] / 1e9) / smem_peak_gbs
        row["persistent"]["smem_roofline"] = "2 x |A| = %.0f KiB per iteration over %d CTAs at 128 B/clk/SM x 1.965 GHz" % (
            smem_bytes / 1024, ctas)
        x, k, out = prob.solve(algo, x0d, mu, {"backend": "persistent", "thres": 1e-8})
        kk, tsec = time_to_tol(out["f_hist"], F_STAR, out["glprox"]["device_ms"] / 1e3)
        row["persistent_thres1e-8"] = {"iters": k, "fval": float(out["fval"]), "ms": out["glprox"]["device_ms"],
                                       "k_to_1e-8_of_fstar": kk, "s_to_1e-8_of_fstar": tsec}
        xr, kr, outr = orc.solve(algo, x0, A, b, mu, {})
        kk, tsec = time_to_tol(outr["f_hist"], float(outr["fval"]), outr["tt"])
        row["cpu_port"] = {"iters": kr, "fval": float(outr["fval"]), "ms": 1e3 * outr["tt"], "it_per_s": kr / outr["tt"],
                           "k_to_1e-8_of_final": kk, "s_to_1e-8_of_final": tsec, "cores": blas_threads()}
        row["rel_obj_diff_gpu_vs_cpu"] = abs(row["persistent"]["fval"] - row["cpu_port"]["fval"]) / abs(row["cpu_port"]["fval"])
        row["published_cpu_seconds"] = {"FProxGD": 1.09, "ProxGD": 1.53}[algo]   # doc/report.md:446-447
        res[algo] = row
    # the other solvers that run on the same kernels, persistent / stream backends
    extra = {}
    for algo, published in (("FGD", 1.24), ("GD", 2.41), ("SGD", 2.02)):   # doc/report.md:443-445
        best = None
        for _ in range(2):
            x, k, out = prob.solve(algo, x0d, mu, {})
            best = out["glprox"]["device_ms"] if best is None else min(best, out["glprox"]["device_ms"])
        extra[algo] = {"iters": k, "fval": float(out["fval"]), "ms": best, "backend": out["glprox"]["backend"],
                       "published_cpu_seconds": published}
    best = None
    for _ in range(2):
        x, k, out = prob.solve_admm_dual(x0d, mu, {})
        best = out["glprox"]["device_ms"] if best is None else min(best, out["glprox"]["device_ms"])
    xr, kr, outr = orc.solve_admm_dual(x0, A, b, mu, {})
    extra["ADMM_dual"] = {"iters": k, "fval": float(out["fval"]), "ms": best, "launches": out["glprox"]["n_launches"],
                          "cpu_port": {"iters": kr, "fval": float(outr["fval"]), "ms": 1e3 * outr["tt"], "cores": blas_threads()},
                          "rel_obj_diff_gpu_vs_cpu": abs(float(out["fval"]) - float(outr["fval"])) / abs(float(outr["fval"])),
                          "published_cpu_seconds": 2.12}   # doc/report.md:571
    res["other_solvers"] = extra
    prob.close()
    return res


# ---- CPU legs -----------------------------------------------------------------------------------------
def blas_threads():
    try:
        from threadpoolctl import threadpool_info
        return max([p.get("num_threads", 1) for p in threadpool_info()] or [1])
    except Exception:
        return _host_cores()


def cpu_solver(algo):
    """(callable, kind): the unmodified reference function staged under oracle/_ref (kind "reference") when
    oracle/make_ref.py has run, else the oracle port of it (kind "port")."""
    from oracle import make_ref
    staged = make_ref.load()
    if staged is not None:
        return staged[0 if algo == "ProxGD" else 1], "reference"
    from oracle import glprox_oracle as orc
    return (lambda x0, A, b, mu, o: orc.solve(algo, x0, A, b, mu, o)), "port"


def cpu_solve(algo, An, bn, x0n, mu, alpha0, maxit):
    """One CPU solve with every host core: (iterations, seconds, fval, f_hist, kind, threads)."""
    from threadpoolctl import threadpool_limits
    fn, kind = cpu_solver(algo)
    with threadpool_limits(limits=_host_cores()):
        threads = blas_threads()
        t0 = time.perf_counter()
        x, k, out = fn(x0n, An, bn, mu, {"alpha0": alpha0, "maxit": maxit})
        secs = time.perf_counter() - t0
    return int(k), secs, float(out["fval"]), [float(v) for v in out["f_hist"]], kind, threads


def product_library_mapped():
    """True when libglprox.so is mapped into this process (it must not be in the reference arm)."""
    try:
        return any("libglprox.so" in ln for ln in open("/proc/self/maps"))
    except OSError:
        return None


def run_reference(args):
    """--impl reference (see the module docstring).  Never imports glprox / loads libglprox.so."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return None
    from oracle import gen
    es = 4 if args.dtype == "f32" else 8
    need_gb = args.m * args.n * es / 1e9
    rows = args.m
    if mem_available_gb() < 1.3 * need_gb + 4:
        rows = max(64, int(args.m * (mem_available_gb() - 4) / (1.3 * need_gb)) // 64 * 64)
    t0 = time.perf_counter()
    An, bn, x0n, _u, mu = gen.gen_problem(args.m, args.n, args.l, args.seed, np.float32 if args.dtype == "f32" else np.float64,
                                          rows=rows)
    t_gen = time.perf_counter() - t0
    alpha0 = mp_step(args.m, args.n)
    # calibrate the sample: one forward + one transposed product on the rows held, 5 products per FProxGD
    # iteration and one for the exit fval (gl_FProxGD_primal.py:25,63,66 x2,129 / :156)
    from threadpoolctl import threadpool_limits
    with threadpool_limits(limits=_host_cores()):
        An @ x0n
        t0 = time.perf_counter()
        r = An @ x0n
        An.T @ r
        t_pair = time.perf_counter() - t0
    prods = (5 if args.algo == "FProxGD" else 4) * 3 * CPU_SAMPLE_MAXIT + 1
    est_step = 0.5 * t_pair * prods
    n_steps = args.steps + args.warmup
    frac = min(1.0, REFERENCE_BUDGET_S / max(est_step * n_steps, 1e-9))
    use = rows if frac >= 1.0 else max(64, int(rows * frac) // 64 * 64)
    As, bs = An[:use], bn[:use]
    scale = use / args.m    # a row sample does proportionally less work per iteration
    kind = threads = None
    for _ in range(args.warmup):
        cpu_solve(args.algo, As, bs, x0n, mu, alpha0, CPU_SAMPLE_MAXIT)
    iters, secs, fval = 0, 0.0, None
    t0 = time.perf_counter()
    for _ in range(args.steps):
        k, tt, fval, _fh, kind, threads = cpu_solve(args.algo, As, bs, x0n, mu, alpha0, CPU_SAMPLE_MAXIT)
        iters += k
        secs += tt
    wall = time.perf_counter() - t0
    value = iters / secs * scale
    sample = ("%d iterations per step (maxit=%d per stage) on rows 0..%d of %d, all %d columns%s; problem generated on "
              "the host in %.1f s (oracle/glprox_gen.c, bit-identical to the device generator)" % (
                  3 * CPU_SAMPLE_MAXIT, CPU_SAMPLE_MAXIT, use, args.m, args.n,
                  "" if use == args.m else "; it/s scaled by rows/m (cost per iteration is linear in m)", t_gen))
    return {"impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * wall / max(args.steps, 1),
            "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": args.dtype,
            "data": "synthetic", "config": bench_config(args),
            "cpu_baseline": {"value": value, "unit": UNIT, "cores": threads, "kind": kind, "sample": sample},
            "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "sample_final_objective": fval, "host_cores": _host_cores(),
            "omp_num_threads_env": os.environ.get("OMP_NUM_THREADS"), "gpu_launches": 0,
            "product_library_mapped": product_library_mapped()}


def main():
    args = parse_args()
    if args.impl == "reference":
        line = run_reference(args)
        if line is not None:
            print(json.dumps(line), flush=True)
        return
    line, ctx = run_ours(args)
    prob, A, b, x0, mu, opts, rank, world = ctx
    host = None
    try:
        line["parity"] = run_parity_gpu(args, ctx)
    except Exception as e:  # noqa: BLE001  (secondary report; never lose the headline line)
        line["parity"] = {"error": repr(e)}
    if not args.no_converge:
        try:
            line["time_to_tol"] = run_convergence(args, ctx)
        except Exception as e:  # noqa: BLE001
            line["time_to_tol"] = {"error": repr(e)}
    if not args.no_e2e:
        e2e, host = run_e2e(args, ctx, args.steps)
        line["e2e"] = e2e if e2e is not None else {"value": None, "unit": UNIT, "h2d_bytes_per_step": 0,
                                                   "d2h_bytes_per_step": 0, "skipped": host}
        if e2e is None:
            host = None
    if not args.no_extra:
        line["extra"] = run_extra(args, ctx)
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        try:
            line["homework_fixture"] = run_homework(args)
        except Exception as e:  # noqa: BLE001
            line["homework_fixture"] = {"error": repr(e)}
        if host is None:
            need_gb = A.numel() * A.element_size() / 1e9
            if mem_available_gb() >= 1.6 * need_gb + 4:
                host = (A.cpu().numpy(), b.cpu().numpy(), x0.cpu().numpy())
        if host is not None:
            # the CPU sample: same inputs (downloaded from HBM), same options, CPU_SAMPLE_MAXIT iterations per stage;
            # it is both the reported cpu_baseline and the checker of `parity`
            An, bn, x0n = host
            k, tt, fval, fh, kind, threads = cpu_solve(args.algo, An, bn, x0n, mu, opts["alpha0"], CPU_SAMPLE_MAXIT)
            line["cpu_baseline"] = {"value": k / tt, "unit": UNIT, "cores": threads, "kind": kind,
                                    "sample": "%d iterations (maxit=%d per stage) of the full-size workload on the inputs "
                                              "downloaded from HBM; %s on NumPy/BLAS" % (
                                                  k, CPU_SAMPLE_MAXIT, "unmodified reference solver (oracle/_ref)"
                                                  if kind == "reference" else "oracle port of the reference"),
                                    "seconds": tt}
            par = line.get("parity", {})
            if "gpu_fval" in par:
                gh = np.asarray(par["gpu_f_hist"])
                ch = np.asarray(fh)
                par["cpu_fval"] = fval
                par["cpu_arithmetic"] = "NumPy %s (what the reference computes on these arrays)" % An.dtype
                par["rel_diff"] = abs(par["gpu_fval"] - fval) / abs(fval)
                par["iters_match"] = bool(par["iters"] == k)
                par["f_hist_max_rel_diff"] = (float(np.max(np.abs(gh - ch) / np.abs(ch)))
                                              if gh.shape == ch.shape and ch.size else None)
                par["ok"] = bool(par["rel_diff"] <= par["bar"] and par["iters_match"])
        else:
            line["cpu_baseline"] = {"value": None, "unit": UNIT, "cores": blas_threads(), "kind": "port",
                                    "sample": "skipped: host RAM too small"}
    if "parity" in line and isinstance(line["parity"], dict):
        line["parity"].pop("gpu_f_hist", None)
    if world > 1:
        import torch.distributed as dist
        dist.barrier()
    if rank == 0:
        sys.stderr.flush()
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
