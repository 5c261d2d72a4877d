#!/usr/bin/env python
"""bench.py -- RAILS outer iterations/s on B200 (BASELINE.json metric).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference]
                    [--workload C1|C2|C3|C4|C5] [--grid G]

A "step" is one outer RAILS iteration (src/LyapunovSolver.hpp:136-343:
projection update, projected solve, L-step residual Lanczos, convergence test,
expansion + CGS2, or a restart compression).  The workload at N = 1 is config
C3 of BASELINE.json -- the 3-D 7-point Laplacian 256^3 (m = 16 777 216), lumped
mass matrix M, B random m x 1, tolerance 1e-4 -- the configuration the
north_star quotes its targets on; it fits one GPU.  (`--workload C3I` is the
same problem with M = I, the equation the C++ reference solves: it never reads
M)  With N > 1 the SAME matrix is row-partitioned over the ranks
(Epetra linear map): "scaling": "strong".

Timed region (`value`): iterations W+1 .. W+K of one solve started from scratch
with the convergence test off (so exactly K iterations run for any K), after
one untimed solve of the same kind that absorbs the process's one-off costs
(memory-pool growth, lazy kernel loading; reported as first_solve_wall_s),
bracketed by CUDA events on the library stream with everything resident in HBM
(each rank reports its own event time; the max over ranks is used).  Inputs are
far larger than the 126 MB L2 (V alone is 134 MB per column), so there is no
L2 flush between iterations.
The window is run --windows times back to back (default 5): `value` is the MEDIAN
window, all of them are listed in `window_ms_per_step` with their spread.  (Round 1
reported the fastest of 3 and hid windows of 18-99 ms; the cause, the stream-ordered
memory pool, is gone: profiles/r02_window_variability.md.)  The reference arm runs
ONE window; the median-of-5 of this arm is the conservative side of that asymmetry.
`e2e`: the solve to the configuration's tolerance through the user-facing C
entry point rails_host_solve_m with HOST buffers -- the CSR arrays of A (and M)
and B are uploaded, V and T downloaded, inside the timed wall-clock region;
value = iterations to tolerance / that time (also reported as
time_to_tolerance_s).
`e2e_tsqr`: the same end-to-end solve with "Residual method" = 2 (exact
residual eigenpairs from the tall-skinny QR of [B, V, AV, MV]; projection,
Gram-Schmidt and restarts on the R factors).  A different sequence of iterates
from the reference's Lanczos estimate, so it is reported beside `e2e`, not as it.
`roofline`: the dominant kernel (first kernel of a step of the device-resident
residual-Lanczos chain: fused update + [V AV B]'w) timed alone with CUDA events
at the workload's row count.
`cpu_baseline` / `--impl reference`: the CPU oracle (CSR + OpenMP restatement
of the reference loop; the reference's own sparse path needs Trilinos, which
is not installable here) on the same inputs with all host threads.
"""
import argparse
import ctypes
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
for p in (ROOT, os.path.join(ROOT, "tests"), os.path.join(ROOT, "oracle")):
    if p not in sys.path:
        sys.path.insert(0, p)

from rails_b200 import problems  # noqa: E402


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default=os.environ.get("RAILS_BENCH_WORKLOAD", "C3"))
    ap.add_argument("--grid", type=int, default=int(os.environ.get("RAILS_BENCH_GRID", "0")),
                    help="override the grid size g of the workload (testing)")
    ap.add_argument("--windows", type=int, default=5,
                    help="identical timed windows run back to back; the MEDIAN is reported, all are listed")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-kernels", action="store_true")
    ap.add_argument("--parity-full", action="store_true",
                    help="solve the workload to tolerance on the GPU and on the CPU arm; print the comparison")
    ap.add_argument("--parity-cpu-grid", type=int, default=0,
                    help="with --parity-full: grid size of the CPU comparison (default: full size)")
    ap.add_argument("--continuation", action="store_true",
                    help="add the two-step continuation leg (default for --workload C5)")
    ap.add_argument("--no-continuation", action="store_true")
    ap.add_argument("--kernels-only", action="store_true",
                    help="only the per-kernel timings (the command profiled with ncu --set full)")
    ap.add_argument("--kernel-k", type=int, default=64)
    ap.add_argument("--step-k", type=int, default=0,
                    help="space size the step kernel is timed at (0: median of the timed window)")
    ap.add_argument("--kernel-reps", type=int, default=10)
    ap.add_argument("--cpu-seconds", type=float, default=25.0,
                    help="time budget of the bounded CPU sample")
    return ap.parse_args()


def workload(args):
    kind, g, dim, n_b, params = problems.CONFIGS[args.workload]
    if args.grid:
        g = args.grid
    params = dict(params)
    return kind, g, dim, n_b, params


def workload_name(args, kind, g, dim, n_b, params=None):
    mass = "lumped mass M" if params and params.get("Use mass matrix") else "M=I"
    return "%s %s %dD g=%d (m=%d), %s, B random m x %d, tol 1e-4" % (
        args.workload, kind, dim, g, g ** dim, mass, n_b)


def window_params(params, warmup, steps):
    """The timed window is iterations W+1..W+K of ONE solve.  The convergence
    test is switched off for it (tolerance 0), so that exactly K iterations run
    whatever K the caller asks for; the solve to the real tolerance is the
    `e2e` leg."""
    P = dict(params)
    P["Tolerance"] = 0.0
    P["Maximum iterations"] = warmup + steps
    return P


class ClockSampler:
    """nvidia-smi clocks / throttle reasons during the timed region, one sample per
    second (round 1 polled every 200 ms; measured in round 2: the poller makes no
    difference to the windows, profiles/r02_window_variability.md -- the spread came
    from the stream-ordered memory pool)."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,"
         "clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index=0):
        self.index = index
        self.proc = None
        self.lines = []

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q,
                 "--format=csv,noheader,nounits", "-lms", "1000"],
                stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.lines.append((time.time(), line.strip()))

    def stop(self, t0=None, t1=None):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.25)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        for ts, line in self.lines:
            if t0 is not None and (ts < t0 or ts > t1 + 0.3):
                continue
            f = [x.strip() for x in line.split(",")]
            if len(f) < 9:
                continue
            try:
                sm.append(float(f[1])); mx.append(float(f[2]))
            except ValueError:
                continue
            for name, val in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown",
                                  "sw_power_cap"), f[5:9]):
                if val.lower().startswith("active"):
                    reasons.add(name)
        if not sm:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["no samples"]}
        return {"sm_mhz": float(np.median(sm)), "sm_max_mhz": float(max(mx)),
                "reasons": sorted(reasons), "samples": len(sm)}


def time_kernels(api, rows, rp, ci, v, global_rows, row_offset, k=64, reps=10, step_k=0, step_n=1):
    """Per-kernel device time at the workload's row count (CUDA events on the
    library stream, after warm-up; operands are GBs, far beyond L2).
    Algorithmic bytes follow SURVEY.md §8(d) / DESIGN.md."""
    lib = api.cuda()
    P = ctypes.c_void_p
    out = {}

    def timed(fn, reps=reps, warm=3):
        for _ in range(warm):
            fn()
        api.check(lib.rails_cuda_event_record(2), "event")
        for _ in range(reps):
            fn()
        api.check(lib.rails_cuda_event_record(3), "event")
        ms = ctypes.c_float(0)
        api.check(lib.rails_cuda_event_elapsed_ms(2, 3, ctypes.byref(ms)), "elapsed")
        return ms.value / reps

    V = api.MultiVector(rows, k + 2)
    q = api.MultiVector(rows, 2)
    api.check(lib.rails_mv_random(V.h, 0, k + 2, row_offset, ctypes.c_uint(7)), "random")
    api.check(lib.rails_mv_random(q.h, 0, 2, row_offset, ctypes.c_uint(8)), "random")
    small = api.MultiVector(k, 1, replicated=True)
    coef = np.asfortranarray(np.full((k, 1), 1e-3))
    m = rows
    nnz = int(rp[-1])

    def rec(name, ms, nbytes, launches):
        out[name] = {"ms": ms, "bytes": nbytes, "gbs": nbytes / ms / 1e6, "launches_per_call": launches}

    ms = timed(lambda: api.check(lib.rails_mv_dot_dev(V.h, 0, k, q.h, 0, 1, small.h, 0), "dot"))
    rec("dot_kx1", ms, 8.0 * m * (k + 1), 2)
    ms = timed(lambda: api.check(lib.rails_mv_gemm(V.h, 0, k, api.dptr(coef), k, 1, 1.0, 1.0,
                                                   q.h, 1), "gemm"))
    rec("gemm_kx1_acc", ms, 8.0 * m * (k + 1) + 8.0 * m, 1)
    A = P()
    api.check(lib.rails_csr_create(rows, global_rows, global_rows, row_offset, api.iptr(rp),
                                   api.iptr(ci), api.dptr(v), ctypes.byref(A)), "csr_create")
    ms = timed(lambda: api.check(lib.rails_csr_spmm(A, 0, q.h, 0, 1, V.h, k), "spmm"))
    rec("spmm_e1", ms, 12.0 * nnz + 4.0 * (m + 1) + 16.0 * m, 1)
    lib.rails_csr_free(A)
    ms = timed(lambda: api.check(lib.rails_mv_axpy(q.h, 1, 0.5, q.h, 0, 1), "axpy"))
    rec("axpy", ms, 24.0 * m, 1)
    # CGS2 of column k against the k columns before it (they need not be
    # orthonormal for the traffic to be the same)
    ms = timed(lambda: api.check(lib.rails_mv_orthogonalize(V.h, 0, k, k), "cgs2"), reps=4, warm=1)
    rec("cgs2_1_vs_k", ms, 32.0 * m * k + 128.0 * m, 12)
    # fused Lanczos step kernels on [V(:, :k/2), V(:, k/2:k), q1] (2 * (k/2) + 1 columns)
    h = k // 2
    hs = (P * 3)(V.h, V.h, q.h)
    c0 = (ctypes.c_int * 3)(0, h, 1)
    nn = (ctypes.c_int * 3)(h, h, 1)
    g = np.zeros(2 * h + 1)
    cf = np.full(2 * h + 1, 1e-3)
    al = ctypes.c_double(0)
    ms = timed(lambda: api.check(lib.rails_mv_multi_dot(3, hs, c0, nn, q.h, 0, 1.0, api.dptr(g)), "multi_dot"))
    rec("multi_dot", ms, 8.0 * m * (2 * h + 2), 2)
    ms = timed(lambda: api.check(lib.rails_mv_multi_dot(3, hs, c0, nn, q.h, 0, 1.0000001, api.dptr(g)), "multi_dot"))
    rec("multi_dot_scale_q", ms, 8.0 * m * (2 * h + 3), 2)
    ms = timed(lambda: api.check(lib.rails_mv_multi_gemv(3, hs, c0, nn, api.dptr(cf), V.h, k, q.h, 0,
                                                         ctypes.byref(al)), "multi_gemv"))
    rec("multi_gemv_alpha", ms, 8.0 * m * (2 * h + 3), 2)
    ms = timed(lambda: api.check(lib.rails_mv_lanczos_update(V.h, k, 1e-3, q.h, 0, 1e-3, q.h, 1,
                                                             ctypes.byref(al)), "update"))
    rec("lanczos_update", ms, 32.0 * m, 2)
    # first kernel of a step of the device-resident chain: fused update + projection
    # (reads raw, prev, prev2 and the 2h+1 block columns, writes w and the normalised prev)
    g1 = np.zeros(2 * h + 2)
    ms = timed(lambda: api.check(lib.rails_mv_lanczos_project(3, hs, c0, nn, V.h, k, k + 1, -1, 1e-3,
                                                              api.dptr(g1)), "project"))
    # + 8m: the squared norm of `prev` is recomputed by this entry point (in the
    # chain it is already on the device)
    rec("lanczos_project", ms, 8.0 * m * (2 * h + 1 + 5), 4)
    # The dominant kernel of an iteration since round 2: one Lanczos step as ONE kernel
    # (lanczos_step_kernel: the row tile of [V AV B] staged once in shared memory by TMA;
    # gemv with coef, three-term update and the next projection out of the staged tile).
    # Timed alone by events around each of its launches inside rails_mv_lanczos_run
    # (rails_lanczos_step_timing), at `step_k` columns per block (the median space size
    # of the timed window).  Algorithmic bytes, each operand read once and each result
    # written once: 8 m (2k + n) for the blocks, q_{i-1} and w_i read, q_i and w_{i+1}
    # written = 8 m (2k + n + 4).
    if step_k:
        sk, n1, Ls = step_k, step_n, 10
        Vb, AVb, Bb = api.MultiVector(rows, sk), api.MultiVector(rows, sk), api.MultiVector(rows, n1)
        Qb = api.MultiVector(rows, Ls + 1)
        for mv, c, sd in ((Vb, sk, 3), (AVb, sk, 4), (Bb, n1, 5)):
            api.check(lib.rails_mv_random(mv.h, 0, c, row_offset, ctypes.c_uint(sd)), "random")
            api.check(lib.rails_mv_scal(mv.h, 0, c, 1.0 / np.sqrt(global_rows)), "scal")
        Tm = np.asfortranarray(-1e-2 * np.eye(sk))
        hb = (P * 3)(Vb.h, AVb.h, Bb.h)
        cb = (ctypes.c_int * 3)(0, 0, 0)
        nb = (ctypes.c_int * 3)(sk, sk, n1)
        alb, beb = np.zeros(Ls), np.zeros(Ls)
        stp = ctypes.c_int(0)

        def chain():
            api.check(lib.rails_mv_random(Qb.h, 0, 1, row_offset, ctypes.c_uint(9)), "random")
            api.check(lib.rails_mv_lanczos_run(3, hb, cb, nb, api.dptr(Tm), sk, Qb.h, 0, Ls, api.dptr(alb),
                                               api.dptr(beb), ctypes.byref(stp)), "lanczos_run")

        for _ in range(3):
            chain()
        api.check(lib.rails_lanczos_step_timing(1), "step_timing")
        for _ in range(max(reps // 2, 3)):
            chain()
        tot, nl = ctypes.c_double(0), ctypes.c_int(0)
        shp = (ctypes.c_int * 5)()
        api.check(lib.rails_lanczos_step_timing_read(ctypes.byref(tot), ctypes.byref(nl), shp), "step_timing_read")
        api.check(lib.rails_lanczos_step_timing(0), "step_timing")
        if nl.value > 0:
            rec("lanczos_step", tot.value / nl.value, 8.0 * m * (2 * sk + n1 + 4), 1)
            out["lanczos_step"].update({"block_columns": int(shp[0]), "tile_rows": int(shp[1]), "stages": int(shp[2]),
                                        "column_groups": int(shp[3]), "ctas": int(shp[4]), "launches_timed": nl.value,
                                        "space_size": sk, "survey_two_pass_bound_bytes": 16.0 * m * (2 * sk + n1 + 2)})
        for mv in (Vb, AVb, Bb, Qb):
            mv.free()
    # The two calls of the path that are bound by FP64 arithmetic rather than by HBM
    # once k, r >~ 50: the restart compression V <- V X (k x r,
    # src/LyapunovSolver.hpp:265-297) and the warm-start projection V'(AV) (k x k,
    # :173-195).  Reported against both ceilings; the FP64 one is measured here with
    # register-only kernels (rails_cuda_fp64_peak: DFMA and DMMA, the larger counts),
    # next to the nominal 37 TFLOP/s of an HGX B200 GPU.
    FP64_NOMINAL_TFLOPS = 37.0
    peak_fma, peak_mma = ctypes.c_double(0), ctypes.c_double(0)
    api.check(lib.rails_cuda_fp64_peak(0, ctypes.byref(peak_fma)), "fp64_peak")
    api.check(lib.rails_cuda_fp64_peak(1, ctypes.byref(peak_mma)), "fp64_peak")
    fp64_measured = max(peak_fma.value, peak_mma.value)
    r = max(k // 2, 1)
    Out = api.MultiVector(rows, r)
    Xc = np.asfortranarray(np.random.RandomState(3).uniform(-1, 1, (k, r)) * 1e-2)
    G = np.zeros((k, k), order="F")

    def rec_flops(name, fn, nbytes, flops):
        l0 = lib.rails_cuda_kernel_launches()
        fn()
        launches = int(lib.rails_cuda_kernel_launches() - l0)
        ms = timed(fn, reps=4, warm=1)
        rec(name, ms, nbytes, launches)
        out[name].update({"flops": flops, "tflops": flops / ms / 1e9,
                          "frac_of_fp64_measured": flops / ms / 1e9 / fp64_measured,
                          "fp64_measured_tflops": {"dfma": peak_fma.value, "dmma": peak_mma.value},
                          "frac_of_fp64_nominal": flops / ms / 1e9 / FP64_NOMINAL_TFLOPS,
                          "fp64_nominal_tflops": FP64_NOMINAL_TFLOPS})

    rec_flops("gemm_kxr_restart",
              lambda: api.check(lib.rails_mv_gemm(V.h, 0, k, api.dptr(Xc), k, r, 1.0, 0.0, Out.h, 0), "gemm"),
              8.0 * m * (k + r), 2.0 * m * k * r)
    rec_flops("dot_kxk_projection",
              lambda: api.check(lib.rails_mv_dot(V.h, 0, k, V.h, 0, k, api.dptr(G), k), "dot"),
              8.0 * m * k, 2.0 * m * k * k)
    Out.free()
    for o in out.values():
        o["k"] = k
    V.free(); q.free(); small.free()
    return out


def continuation_leg(api, lib, dist, world, rp, ci, v, n_b, params, B, m, r0, M):
    """first: solve with A.  warm: A <- A - 0.01 I in place, solve from the basis of
    `first` (resident).  cold: the same second problem from a random start.  Wall
    clock per solve between barriers, max over ranks; X_warm against X_cold through
    the k x k Gram matrices (all-reduced)."""
    def barrier():
        lib.rails_cuda_sync()
        if dist is not None:
            dist.barrier()

    def timed(fn):
        barrier()
        t = time.time()
        out = fn()
        barrier()
        return out, allreduce_max(dist, time.time() - t)

    def gram(X, Y):
        G = np.ascontiguousarray(X.T @ Y)
        if dist is not None:
            import torch
            t = torch.from_numpy(G)
            dist.all_reduce(t)
            G = t.numpy()
        return G

    ses, t_setup = timed(lambda: api.Session(rp, ci, v, n_b, B=B, global_rows=m, row_offset=r0, M=M, skip_draws=1))
    P = dict(params)
    P["Random seed"] = 12345          # every solve of this leg draws the same start vectors
    first, t_first = timed(lambda: ses.solve(P))
    v2 = problems.shift_diagonal(rp, ci, v, -0.01, r0)
    _, t_update = timed(lambda: ses.update_operator(v2))
    Pw = dict(P)
    Pw["Restart from solution"] = 1
    warm, t_warm = timed(lambda: ses.solve(Pw))
    Vw, Tw = ses.download()
    cold, t_cold2 = timed(lambda: ses.solve(P))
    Vc, Tc = ses.download()
    ses.free()
    Gww, Gcc, Gwc = gram(Vw, Vw), gram(Vc, Vc), gram(Vw, Vc)
    nw = np.trace(Tw @ Gww @ Tw.T @ Gww)
    nc = np.trace(Tc @ Gcc @ Tc.T @ Gcc)
    x_diff = float(np.sqrt(max(nw - 2 * np.trace(Tw @ Gwc @ Tc.T @ Gwc.T) + nc, 0.0) / nc))

    def rec(o, t):
        return {"ret": int(o["ret"]), "iterations": int(o["iterations"]), "space_size": int(o["k"]),
                "wall_s": t, "solve_s": o["solve_seconds"], "final_relative_residual": float(o["rel"][-1]),
                "first_relative_residual": float(o["rel"][0])}

    return {"what": "A, then A - 0.01 I: warm = \"Restart from solution\" from the resident basis of the first "
                    "solve (operator values replaced in place), cold = the same problem from a random start",
            "session_setup_s": t_setup, "operator_update_s": t_update,
            "first": rec(first, t_first), "warm": rec(warm, t_warm), "cold": rec(cold, t_cold2),
            "warm_over_cold_wall": t_warm / t_cold2, "x_rel_diff_warm_vs_cold": x_diff}


def parity_full(args):
    """`--parity-full`: the workload solved TO TOLERANCE at full size on the GPU (through
    rails_host_solve, host buffers) and by the CPU arm (the oracle port, all host
    threads; the unmodified reference loop is ~20x slower than the port and is compared
    at reduced sizes in tests/), same inputs: iteration counts, space sizes, the largest
    relative difference of the residual histories and of X.  `--parity-cpu-grid G` runs
    the CPU side (and a second GPU solve) at grid G instead, for workloads whose CPU
    solve does not fit a bench run (C4: ~1 min per CPU iteration at full size).
    One JSON line; not a bench line."""
    import harness
    from rails_b200 import api
    api.init(0)
    api.host()
    kind, g, dim, n_b, params = workload(args)
    out = {"parity_full": workload_name(args, kind, g, dim, n_b, params), "params": params}

    def both(grid, cpu):
        m = grid ** dim
        rp, ci, v = problems.build(kind, grid, dim)
        M = problems.lumped_mass(m) if params.get("Use mass matrix") else None
        B = problems.reference_B(m, n_b)
        P = dict(params)
        P["Skip random draws"] = 1
        cap = int(P.get("Restart size", 0)) if P.get("Restart size", 0) > 0 else 400
        cap = max(cap + int(P.get("Expand size", 1)), 64)
        t = time.time()
        d = api.solve(rp, ci, v, n_b, P, B=B, v_cap_cols=cap, M=M)
        rec = {"grid": grid, "rows": m,
               "gpu": {"ret": int(d["ret"]), "iterations": int(d["iterations"]), "space_size": int(d["k"]),
                       "max_space_size": int(d["space"].max()), "wall_s": time.time() - t,
                       "final_relative_residual": float(d["rel"][-1])}}
        if not cpu:
            return rec
        harness.oracle().oracle_set_num_threads(os.cpu_count() or 1)
        t = time.time()
        o = harness.oracle_solve(rp, ci, v, n_b, P, B=B, M=M, v_cap_cols=cap)
        rec["cpu"] = {"kind": "port", "cores": harness.oracle().oracle_num_threads(), "ret": int(o["ret"]),
                      "iterations": int(o["iterations"]), "space_size": int(o["k"]),
                      "max_space_size": int(o["space"].max()), "wall_s": time.time() - t,
                      "final_relative_residual": float(o["rel"][-1])}
        n = min(d["iterations"], o["iterations"])
        rel = np.abs(d["rel"][:n] - o["rel"][:n]) / o["rel"][:n]
        rec["history_rel_diff_first10"] = float(rel[:10].max())
        rec["history_rel_diff_all_common"] = float(rel.max())
        rec["first_iteration_above_1e-8"] = int(np.argmax(rel > 1e-8)) + 1 if np.any(rel > 1e-8) else None
        rec["space_sizes_equal_on_common_prefix"] = bool(np.array_equal(d["space"][:n], o["space"][:n]))
        if "V" in d and "V" in o:
            rec["x_rel_diff"] = float(harness.lowrank_rel_diff(d["V"], d["T"], o["V"], o["T"]))
        return rec

    if args.parity_cpu_grid:
        out["full_size"] = both(g, cpu=False)
        out["reduced"] = both(args.parity_cpu_grid, cpu=True)
    else:
        out["full_size"] = both(g, cpu=True)
    print(json.dumps(out), flush=True)


def dist_setup(n_gpus):
    """torch.distributed (gloo) is the control plane only: rendezvous, barrier,
    max-over-ranks of the device times, broadcast of the NCCL id.  The data
    path (halo exchange, all-reduce of the k x e blocks) is NCCL inside
    librails_cuda.so."""
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    dist = None
    if world > 1:
        import torch.distributed as dist_mod
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist_mod.init_process_group(backend="gloo", rank=rank, world_size=world)
        dist = dist_mod
    return dist, rank, world, local


def allreduce_max(dist, x):
    if dist is None:
        return x
    import torch
    t = torch.tensor([x], dtype=torch.float64)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t[0])


def cpu_sample(rp, ci, v, n_b, params, B, warmup, steps, budget_s, M=None):
    """The CPU oracle on the same inputs, all host threads; iterations
    warmup+1 .. warmup+steps, cut short after `budget_s` seconds."""
    import harness
    P = window_params(params, warmup, steps)
    P["Oracle time limit"] = budget_s
    harness.oracle().oracle_set_num_threads(os.cpu_count() or 1)      # all host threads
    o = harness.oracle_solve(rp, ci, v, n_b, P, B=B, M=M)
    sec = o["seconds"]
    done = len(sec)
    cores = harness.oracle().oracle_num_threads()
    if done > warmup + 1:
        n = done - warmup
        rate = n / (sec[-1] - sec[warmup - 1]) if warmup > 0 else n / sec[-1]
        sample = "iterations %d..%d of the same solve, full size" % (warmup + 1, done)
    else:
        rate = done / sec[-1]
        sample = "iterations 1..%d of the same solve, full size (budget hit inside warm-up)" % done
    return {"value": rate, "unit": "iterations/s", "cores": cores, "kind": "port",
            "sample": sample, "iterations_run": done, "seconds": float(sec[-1]),
            "rel_residual_history": [float(x) for x in o["rel"][:done]]}


def ref_loop_sample(rp, ci, v, n_b, params, B, warmup, steps, budget_s):
    """The UNMODIFIED reference loop (src/LyapunovSolver.hpp) over the reference's own
    StlWrapper and a CSR/OpenMP operator (oracle/_ref/librails_ref_csr.so), all host
    threads, on the same operator and B with M = I (the C++ loop never reads M).  The
    loop cannot be stopped from outside, so a 3-iteration run sizes the sample to the
    time budget; the rate is taken between the progress lines of the iterations."""
    import harness
    if harness.ref_csr() is None:
        return None
    P = window_params(params, warmup, steps)
    P.pop("Use mass matrix", None)
    P.pop("Skip random draws", None)
    probe = dict(P)
    probe["Maximum iterations"] = 3
    t0 = time.time()
    harness.ref_csr_solve(rp, ci, v, n_b, probe, B=B, skip_draws=1)
    per_it = (time.time() - t0) / 3.0
    # iterations get dearer as the space grows: leave a factor 2
    n = int(max(4, min(warmup + steps, (budget_s - 3 * per_it) / (2.0 * per_it))))
    P["Maximum iterations"] = n
    r = harness.ref_csr_solve(rp, ci, v, n_b, P, B=B, skip_draws=1)
    sec = r["seconds"]
    done = len(sec)
    first = min(warmup, max(done - 2, 0))
    rate = (done - first) / (sec[-1] - (sec[first - 1] if first > 0 else 0.0))
    return {"value": rate, "unit": "iterations/s", "cores": os.cpu_count() or 1, "kind": "reference",
            "what": "unmodified reference loop + reference StlWrapper + CSR/OpenMP operator "
                    "(oracle/_ref/librails_ref_csr.so), M = I",
            "sample": "iterations %d..%d of the same solve, full size" % (first + 1, done),
            "iterations_run": done, "seconds": float(sec[-1]),
            "rel_residual_history": [float(x) for x in r["rel"][:done]]}


def run_reference(args):
    """--impl reference: the reference's CPU implementation of the path.  The
    reference's sparse CPU backend (Epetra) cannot be built here, so this is
    the oracle port (CSR + OpenMP wrappers' arithmetic under the reference's
    loop), with all host threads; rank 0 only."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    kind, g, dim, n_b, params = workload(args)
    rp, ci, v = problems.build(kind, g, dim)
    M = problems.lumped_mass(g ** dim) if params.get("Use mass matrix") else None
    # `std::srand(1); B.random()` of the reference, made on the host
    B = problems.reference_B(g ** dim, n_b)
    params["Skip random draws"] = 1
    budget = float(os.environ.get("RAILS_BENCH_REF_SECONDS", "150"))
    t0 = time.time()
    extra = {}
    if M is None:
        # M = I: the reference's own loop solves this workload -- it IS the arm
        cb = ref_loop_sample(rp, ci, v, n_b, params, B, args.warmup, args.steps, budget)
        if cb is None:
            cb = cpu_sample(rp, ci, v, n_b, params, B, args.warmup, args.steps, budget, M=M)
    else:
        # generalised equation (lumped M): the C++ reference never reads M, so
        # the arm is the CPU restatement of the MATLAB branch; the reference loop on the
        # same operator and B with M = I is timed beside it (same kernels per iteration
        # minus the M products)
        cb = cpu_sample(rp, ci, v, n_b, params, B, args.warmup, args.steps, 0.6 * budget, M=M)
        # The reference loop deep-copies full-capacity multivectors on every std::move
        # (~20x the port's time per iteration), so above ~2 M rows even its
        # 3-iteration probe exceeds the budget: it is then timed on the same stencil at a
        # REDUCED grid, stated in the record and never extrapolated.
        g_side = g
        while g_side ** dim > 2200000:
            g_side //= 2
        if g_side == g:
            rl = ref_loop_sample(rp, ci, v, n_b, params, B, args.warmup, args.steps, 0.4 * budget)
        else:
            rp2, ci2, v2 = problems.build(kind, g_side, dim)
            rl = ref_loop_sample(rp2, ci2, v2, n_b, params, problems.reference_B(g_side ** dim, n_b),
                                 args.warmup, args.steps, 0.4 * budget)
        if rl is not None:
            extra["reference_loop_M_is_I"] = {k: rl[k] for k in ("value", "unit", "cores", "kind", "what", "sample")}
            extra["reference_loop_M_is_I"]["rows"] = g_side ** dim
            extra["reference_loop_M_is_I"]["grid"] = g_side
            if g_side != g:
                extra["reference_loop_M_is_I"]["sample"] = extra["reference_loop_M_is_I"]["sample"].replace(
                    "full size", "REDUCED size g=%d (m=%d)" % (g_side, g_side ** dim))
    line = {
        "impl": "reference", "metric": "RAILS outer iterations/s", "value": cb["value"],
        "unit": "iterations/s", "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": 1e3 / cb["value"], "higher_is_better": True, "scaling": "strong",
        "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": workload_name(args, kind, g, dim, n_b, params), "params": params,
                   "window": "iterations W+1..W+K of one solve, convergence test off; ONE window "
                             "(the b200 arm reports the median of its windows)"},
        "cpu_baseline": {k: cb[k] for k in ("value", "unit", "cores", "kind", "sample")},
        "e2e": {"value": cb["value"], "unit": "iterations/s", "h2d_bytes_per_step": 0,
                "d2h_bytes_per_step": 0},
        "gpu_launches": 0, "wall_s": time.time() - t0,
    }
    line.update(extra)
    print(json.dumps(line), flush=True)


def run_b200(args):
    from rails_b200 import api
    dist, rank, world, local = dist_setup(args.gpus)
    if world != args.gpus:
        raise SystemExit("--gpus %d but WORLD_SIZE=%d (launch N>1 with torch.distributed.run)"
                         % (args.gpus, world))
    api.init(local)
    lib = api.cuda()
    api.host()
    if world > 1:
        ident = [None]
        if rank == 0:
            buf = ctypes.create_string_buffer(128)
            api.check(lib.rails_cuda_comm_unique_id(buf), "comm_unique_id")
            ident[0] = buf.raw
        dist.broadcast_object_list(ident, src=0)
        api.check(lib.rails_cuda_comm_init(rank, world, ident[0]), "comm_init")

    kind, g, dim, n_b, params = workload(args)
    m = g ** dim
    r0, r1 = problems.partition_rows(m, world, rank)
    rows = r1 - r0
    rp, ci, v = problems.build(kind, g, dim, r0, r1)
    M = problems.lumped_mass(m, r0, r1) if params.get("Use mass matrix") else None
    # B on the host (so that it is an uploaded input), same for every arm: the
    # reference's own `std::srand(1); B.random()` stream, this
    # rank's rows by jump-ahead; "Skip random draws" keeps the std::rand() sequence
    # of V and of the Lanczos start vectors that of a reference run
    B = problems.reference_B(m, n_b, r0, r1)
    params["Skip random draws"] = 1
    if args.kernels_only:
        kernels = time_kernels(api, rows, rp, ci, v, m, r0, k=args.kernel_k, reps=args.kernel_reps,
                               step_k=args.step_k or 15, step_n=n_b)
        if rank == 0:
            print(json.dumps({"kernels": kernels, "rows": rows}), flush=True)
        return
    W, K = args.warmup, args.steps

    # ---- warm-up solve (untimed): the first solve of a process pays one-off
    #      costs that are not per-iteration work -- growth of the stream-ordered
    #      memory pool (V/AV/Q/restart temporaries, tens of ms per GB), lazy
    #      loading of each kernel on first launch, NCCL/peer-mapping first use.
    #      Measured on C3: first solve 456 ms for the window, second 358 ms.
    t_cold = time.time()
    cold = api.solve(rp, ci, v, n_b, window_params(params, W, K), B=B, global_rows=m, row_offset=r0, M=M)
    lib.rails_cuda_sync()
    t_cold = time.time() - t_cold
    if cold["ret"] >= 100:
        raise SystemExit("rails_host_solve (warm-up) failed with code %d" % cold["ret"])

    # ---- leg 1: device-timed window, everything resident in HBM -------------
    #      The window (W untimed + exactly K timed iterations of one solve) is run
    #      --windows times; the median is reported and all of them are listed.
    #      Every window obeys the timing rules on its own.
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    t0 = time.time()
    all_ms, runs = [], []
    for _ in range(max(1, args.windows)):
        if dist is not None:
            dist.barrier()
        lib.rails_cuda_sync()
        cur = api.solve(rp, ci, v, n_b, window_params(params, W, K), B=B, global_rows=m, row_offset=r0,
                        time_from=W, time_to=W + K, M=M)
        lib.rails_cuda_sync()
        if dist is not None:
            dist.barrier()
        if cur["ret"] >= 100:
            raise SystemExit("rails_host_solve failed with code %d" % cur["ret"])
        if cur["timed_ms"] <= 0.0:
            raise SystemExit("the solve ended after %d iterations, before the timed window "
                             "[%d, %d] closed" % (cur["iterations"], W + 1, W + K))
        cur_ms = allreduce_max(dist, cur["timed_ms"])
        all_ms.append(cur_ms / K)
        runs.append((cur_ms, cur))
    t1 = time.time()
    clocks = sampler.stop(t0, t1) if rank == 0 else None
    # the MEDIAN window is the reported one (lower middle for an even count)
    runs.sort(key=lambda r: r[0])
    timed_ms, res = runs[(len(runs) - 1) // 2]

    # ---- leg 2: end to end through the C entry point, host buffers in and out,
    #      solved to the configuration's tolerance -----------------------------
    cap = 400

    def end_to_end(P, repeats):
        """Median of `repeats` end-to-end solves (host buffers in and out inside the
        timed region); the max over ranks is taken per solve.  Returns the median
        solve, its wall time and all wall times."""
        done = []
        for _ in range(repeats):
            if dist is not None:
                dist.barrier()
            t_start = time.time()
            out = api.solve(rp, ci, v, n_b, P, B=B, global_rows=m, row_offset=r0, v_cap_cols=cap, M=M)
            lib.rails_cuda_sync()
            if dist is not None:
                dist.barrier()
            w = allreduce_max(dist, time.time() - t_start)
            if out["ret"] >= 100:
                raise SystemExit("rails_host_solve (e2e leg) failed with code %d" % out["ret"])
            done.append((w, out))
        walls = [w for w, _ in done]
        done.sort(key=lambda r: r[0])
        w, out = done[(len(done) - 1) // 2]
        return out, w, walls

    e2e, wall, e2e_walls = end_to_end(params, 3)
    h2d = rp.nbytes + ci.nbytes + v.nbytes + B.nbytes + (sum(a.nbytes for a in M) if M else 0)
    d2h = rows * e2e["k"] * 8 + e2e["k"] * e2e["k"] * 8
    h2d_total = allreduce_max(dist, float(h2d)) * world
    d2h_total = allreduce_max(dist, float(d2h)) * world

    # ---- leg 3: the same end-to-end solve with "Residual method" = 2 (exact
    #      residual eigenpairs from the tall-skinny QR of [B, V, AV, MV] and the
    #      whole iteration in the coordinates of that basis, BASELINE.json
    #      north_star item (3)).  Reported beside `e2e`, never instead of it: the
    #      iterates are not the reference's Lanczos iterates. --------------------
    p2 = dict(params)
    p2["Residual method"] = 2
    e2e2, wall2, e2e2_walls = end_to_end(p2, 3)
    x_diff = None
    if world == 1 and "V" in e2e and "V" in e2e2:
        V1, T1, V2, T2 = e2e["V"], e2e["T"], e2e2["V"], e2e2["T"]
        G11, G22, G12 = V1.T @ V1, V2.T @ V2, V1.T @ V2
        n1 = np.trace(T1 @ G11 @ T1.T @ G11)
        n2 = np.trace(T2 @ G22 @ T2.T @ G22)
        x_diff = float(np.sqrt(max(n1 - 2 * np.trace(T1 @ G12 @ T2.T @ G12.T) + n2, 0.0) / n1))

    # ---- leg 4 (workload C5, or --continuation): the two-step continuation of
    #      test/LyapunovSolver_test.cpp:312-352 with everything resident on the
    #      device: one session holds A, B and the solution; the second solve runs
    #      on A - 0.01 I (values replaced in place, same pattern) warm-started from
    #      the basis of the first ("Restart from solution"), and is compared with
    #      the same solve started cold. ------------------------------------------
    continuation = None
    if args.continuation or (args.workload == "C5" and not args.no_continuation):
        continuation = continuation_leg(api, lib, dist, world, rp, ci, v, n_b, params, B, m, r0, M)

    kernels = None
    if not args.no_kernels:
        # the step kernel at the median space size of the timed window
        sp = [int(x) for x in res["space"][W:W + K]] or [15]
        step_k = args.step_k or int(np.median(sp))
        kernels = time_kernels(api, rows, rp, ci, v, m, r0, k=args.kernel_k, reps=args.kernel_reps,
                               step_k=step_k, step_n=n_b)
    if rank != 0:
        return

    import json as _json
    peaks = {}
    try:
        peaks = _json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    peak = float(peaks.get("hbm_gbs", 6650.0))
    peak_src = "measured (MEASURED_PEAKS.json hbm_gbs)" if "hbm_gbs" in peaks else "fallback 6650"
    roofline = None
    if kernels:
        # the dominant kernel of an outer iteration: the one-kernel Lanczos step
        # (L - 1 launches per iteration; ~80 % of the iteration's device time)
        traffic = None
        tj = {}
        try:
            tj = _json.load(open(os.path.join(ROOT, "profiles", "traffic.json")))
        except Exception:
            pass
        if "lanczos_step" in kernels:
            dk = kernels["lanczos_step"]
            # ncu DRAM bytes of one launch at the N = 1 row count and this space size;
            # not a measurement for other shapes or the per-rank blocks of N > 1
            t = tj.get("lanczos_step", {})
            if world == 1 and rows == t.get("rows") and dk["space_size"] == t.get("space_size"):
                traffic = t.get("dram_bytes_per_launch")
            name = ("lanczos_step_kernel (one Lanczos step in one kernel: TMA-staged row tile of [V AV B], "
                    "gemv + update + next projection; %d block columns, space size %d, m=%d rows, "
                    "tile %d rows x %d stages, %d column groups; events around each launch inside "
                    "rails_mv_lanczos_run)" % (dk["block_columns"], dk["space_size"], rows, dk["tile_rows"],
                                               dk["stages"], dk["column_groups"]))
        else:
            dk = kernels["lanczos_project"]
            if world == 1 and rows == 16777216 and dk["k"] == 64:
                traffic = tj.get("lanczos_dot")
            name = ("lanczos_dot_kernel (fused Lanczos update + [V AV B]'w, %d columns, m=%d rows; timed through "
                    "rails_mv_lanczos_project)" % (2 * (dk["k"] // 2) + 1, rows))
        roofline = {"bound": "hbm", "kernel": name,
                    "achieved": dk["gbs"], "peak": peak, "unit": "GB/s", "frac": dk["gbs"] / peak,
                    "traffic": traffic, "peak_source": peak_src,
                    "algorithmic_bytes_per_launch": dk["bytes"], "ms_per_launch": dk["ms"]}

    cpu_baseline = None
    if world == 1 and not args.no_cpu_baseline:
        cb = cpu_sample(rp, ci, v, n_b, params, B, W, K, args.cpu_seconds, M=M)
        cpu_baseline = {k: cb[k] for k in ("value", "unit", "cores", "kind", "sample")}
        n = min(cb["iterations_run"], res["iterations"])
        ref_rel = np.array(cb["rel_residual_history"][:n])
        cpu_baseline["history_rel_diff_vs_gpu"] = float(
            np.max(np.abs(ref_rel - res["rel"][:n]) / ref_rel))

    value = K / (timed_ms / 1e3)
    line = {
        "metric": "RAILS outer iterations/s", "value": value, "unit": "iterations/s",
        "n_gpus": world, "steps": K, "warmup": W, "ms_per_step": timed_ms / K,
        "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f64",
        "data": "synthetic",
        "config": {"workload": workload_name(args, kind, g, dim, n_b, params), "params": params,
                   "window": "iterations W+1..W+K of one solve, convergence test off; median of %d "
                             "identical windows (all listed in window_ms_per_step)" % len(all_ms),
                   "l2": "inputs >> L2 (no flush): V is %.0f MB per column" % (rows * 8 / 1e6),
                   "parallelism": "row partition over %d GPU(s)" % world,
                   "small_reductions": ("peer-memory mailboxes (NVLink P2P)" if lib.rails_cuda_comm_p2p()
                                        else "NCCL all-reduce") if world > 1 else "local",
                   "space_size_in_window": [int(res["space"][W]) if W < len(res["space"]) else None,
                                            int(res["space"][min(W + K, len(res["space"])) - 1])]},
        "e2e": {"value": e2e["iterations"] / wall, "unit": "iterations/s",
                "h2d_bytes_per_step": h2d_total / e2e["iterations"],
                "d2h_bytes_per_step": d2h_total / e2e["iterations"],
                "time_to_tolerance_s": wall, "iterations_to_tolerance": int(e2e["iterations"]),
                "converged": e2e["ret"] == 0, "space_size": int(e2e["k"]),
                "setup_seconds": e2e["setup_seconds"], "solve_seconds": e2e["solve_seconds"],
                "what": "rails_host_solve to tolerance 1e-4 with host CSR/B(/M) in, V/T out; "
                        "%d iterations in %.3f s (median of 3 runs)" % (e2e["iterations"], wall),
                "all_runs_s": e2e_walls},
        "e2e_tsqr": {"value": e2e2["iterations"] / wall2, "unit": "iterations/s",
                     "time_to_tolerance_s": wall2, "iterations_to_tolerance": int(e2e2["iterations"]),
                     "converged": e2e2["ret"] == 0, "space_size": int(e2e2["k"]),
                     "setup_seconds": e2e2["setup_seconds"], "solve_seconds": e2e2["solve_seconds"],
                     "final_relative_residual": float(e2e2["rel"][-1]),
                     "x_rel_diff_vs_e2e": x_diff,
                     "what": "the e2e solve with \"Residual method\" = 2: exact residual eigenpairs "
                             "from the tall-skinny QR of [B, V, AV, MV], iteration carried out on the "
                             "R factors; same tolerance, host buffers in and out (median of 3 runs)",
                     "all_runs_s": e2e2_walls},
        "window_ms_per_step": all_ms,
        "window_ms_per_step_min": min(all_ms), "window_ms_per_step_max": max(all_ms),
        "window_spread": (max(all_ms) - min(all_ms)) / min(all_ms),
        "gpu_launches": int(res["timed_launches"]),
        "host_syncs_per_step": res["timed_syncs"] / float(K),
        "clocks": clocks,
        "roofline": roofline,
        "cpu_baseline": cpu_baseline,
        "kernels": kernels,
        "first_solve_wall_s": t_cold,
        "dense_solve_seconds": res["dense_solve_seconds"],
        "solve_seconds": res["solve_seconds"], "setup_seconds": res["setup_seconds"],
        "relative_residual": [float(e2e["rel"][0]), float(e2e["rel"][-1])],
    }
    if continuation is not None:
        line["continuation"] = continuation
    print(json.dumps(line), flush=True)


def main():
    args = parse_args()
    # The contract is ONE JSON line on stdout.  Native libraries (NCCL's
    # version banner, build recipes) write to file descriptor 1 directly, so
    # fd 1 is pointed at stderr for the duration of the run and the JSON line
    # goes to the saved descriptor.
    sys.stdout.flush()
    saved = os.dup(1)
    os.dup2(2, 1)
    out = os.fdopen(saved, "w")
    real_print = print

    def emit(*a, **kw):
        kw["file"] = out
        real_print(*a, **kw)
        out.flush()
    globals()["print"] = emit
    if args.parity_full:
        parity_full(args)
    elif args.impl == "reference":
        run_reference(args)
    else:
        run_b200(args)


if __name__ == "__main__":
    main()
