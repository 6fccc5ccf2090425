#!/usr/bin/env python
"""Benchmark of the finite_difference time-stepping hot path on B200.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference]
                    [--workload c3|c2|c4|c5] [--scaling weak|strong]
                    [--no-workloads] [--no-history-legs] [--no-cpu-baseline]
                    [--timeline FILE]

One "step" is one time step of the model (every grid point advanced once).
Headline workload: BASELINE.json configs[2], 3-D heat diffusion 512^3 fp64,
7-point stencil, periodic -- the configuration the north-star target is quoted
on.  Prints ONE JSON line (see DESIGN.md "Measurement").

  value      point-updates/s, state resident in HBM, CUDA events on the
             launching stream around exactly K steps driven through the public
             `finite_difference` API (the recorded-statement layer included);
             several ranks: barrier + device synchronisation immediately
             before the first event, maximum over ranks
  e2e        the same K steps through the public API with HOST buffers:
             initial condition uploaded from pinned host memory, final level
             read back to the host, both inside the timed region
  roofline   algorithmic bytes (16 B/point: one read + one write of the state)
             over the measured step time, against the measured HBM copy peak
  workloads  the other BASELINE.json configurations timed the same way in the
             same process after the headline (c2, c4, c5 on one GPU; c5 =
             2048^3 fp32 over the ranks on several), plus the end-to-end cost
             of the reference's keep-every-level history policy
  parity_check  (several ranks) small slab-decomposed models compared bit for
             bit with the numpy oracle before anything is timed; a mismatch
             ends the run with a non-zero exit code
  cpu_baseline  the reference path on this box's host cores (numpy: 1 core):
             the unmodified FD.py when build() staged it (oracle/_ref), else
             the numpy port (oracle/fd_numpy.py)

Clocks are sampled by ONE helper process (started by rank 0 before the timed
regions, NVML, all participating GPUs) so that no benchmark process runs a
sampling thread of its own; samples are matched to the timed regions by their
CLOCK_MONOTONIC time stamps.
"""

from __future__ import annotations

import argparse
import contextlib
import gc
import io
import json
import os
import subprocess
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
for p in (ROOT, os.path.join(ROOT, "finite-difference_b200")):
    if p not in sys.path:
        sys.path.insert(0, p)

WORKLOADS = {
    # name: (description, spatial shape per GPU, bytes per point update)
    "c3": ("3-D heat diffusion 512^3 fp64, 7-point stencil, periodic (BASELINE.json configs[2])", (512, 512, 512), 16),
    "c2": ("2-D heat diffusion 8192^2 fp64, 5-point stencil, fixed BCs (BASELINE.json configs[1])", (8192, 8192), 16),
    "c4": ("2-D wave equation 8192^2 fp64, n_time_ders=2, 9-point-per-axis stencils, periodic (BASELINE.json configs[3])",
           (8192, 8192), 32),
    "c5": ("3-D staggered cell/edge acoustic model fp32, periodic, 256x2048x2048 slab per GPU "
           "(2048^3 on 8 GPUs, BASELINE.json configs[4])", (256, 2048, 2048), 32),
}
DTYPES = {"c5": np.float32}
PERIODIC = {"c3": True, "c2": False, "c4": True, "c5": True}
K_DIFF, DT = 0.1, 1e-3
GRID_SPACING = 1.0     # --dx: 1 (BASELINE synthetic inputs); other values exercise the scaled kernels


# ------------------------------------------------------------------ helpers
def measured_peak():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.isfile(path):
        try:
            return float(json.load(open(path))["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
        except Exception:
            pass
    return 6650.0, "fallback (B200_PROFILING.md)"


def recorded_traffic(workload):
    """dram bytes per launch of the dominant kernel from the committed ncu
    --set full capture (profiles/traffic.json), or None."""
    path = os.path.join(ROOT, "profiles", "traffic.json")
    if os.path.isfile(path):
        try:
            return json.load(open(path)).get(workload, {}).get("dram_bytes_per_launch")
        except Exception:
            return None
    return None


# ------------------------------------------------------------ clock sampling
NVML_REASONS = {"hw_slowdown": 0x8, "hw_thermal_slowdown": 0x40, "sw_thermal_slowdown": 0x20, "sw_power_cap": 0x4}


def clock_helper_main(uuids):
    """Body of the helper process: polls SM clock and clock-event reasons of the
    given GPUs through NVML until stdin closes, then prints every sample
    ([monotonic seconds, gpu, sm MHz, max MHz, reason mask]) as one JSON line."""
    import select
    rows, source = [], None
    try:
        import pynvml as n
        n.nvmlInit()
        handles = []
        for k, u in enumerate(uuids):
            try:
                handles.append(n.nvmlDeviceGetHandleByUUID(u if u.startswith("GPU-") else "GPU-" + u))
            except Exception:
                handles.append(n.nvmlDeviceGetHandleByIndex(k))
        maxes = [n.nvmlDeviceGetMaxClockInfo(h, n.NVML_CLOCK_SM) for h in handles]
        source = "nvml"
    except Exception as exc:            # no NVML binding: nothing to sample with
        print("READY", flush=True)
        sys.stdin.read()
        print(json.dumps({"source": None, "rows": [], "error": repr(exc)}), flush=True)
        return
    print("READY", flush=True)
    while True:
        for k, h in enumerate(handles):
            try:
                sm = n.nvmlDeviceGetClockInfo(h, n.NVML_CLOCK_SM)
                try:
                    mask = n.nvmlDeviceGetCurrentClocksEventReasons(h)
                except Exception:
                    mask = n.nvmlDeviceGetCurrentClocksThrottleReasons(h)
                rows.append([time.monotonic(), k, sm, maxes[k], int(mask)])
            except Exception:
                pass
        if select.select([sys.stdin], [], [], 0.0005)[0]:
            break
        if len(rows) > 2_000_000:
            del rows[::2]
    print(json.dumps({"source": source, "rows": rows}), flush=True)


class ClockHelper:
    """Rank 0's handle on the helper process (None-safe on the other ranks)."""

    def __init__(self, uuids):
        self.proc, self.rows, self.source = None, None, None
        if not uuids:
            return
        try:
            self.proc = subprocess.Popen([sys.executable, os.path.abspath(__file__), "--clock-helper", ",".join(uuids)],
                                         stdin=subprocess.PIPE, stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.proc.stdout.readline()          # "READY": NVML is initialised, sampling has begun
        except OSError:
            self.proc = None

    def finish(self):
        if self.proc is None or self.rows is not None:
            return
        try:
            out, _ = self.proc.communicate(input="", timeout=20)
            got = json.loads(out.strip().splitlines()[-1])
            self.rows, self.source = got["rows"], got["source"]
        except Exception:
            self.rows = []
            with contextlib.suppress(Exception):
                self.proc.kill()

    def summary(self, window):
        """Clocks of all sampled GPUs inside the (t0, t1) window of a timed region."""
        self.finish()
        if not self.rows:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0, "source": self.source}
        t0, t1 = window
        inside = [r for r in self.rows if t0 <= r[0] <= t1]
        if not inside:           # region shorter than one sampling round: the nearest sample of every GPU
            for g in sorted({r[1] for r in self.rows}):
                inside.append(min((r for r in self.rows if r[1] == g), key=lambda r: abs(r[0] - 0.5 * (t0 + t1))))
        sm = [r[2] for r in inside]
        mask = 0
        for r in inside:
            mask |= r[4]
        per_gpu = {}
        for g in sorted({r[1] for r in inside}):
            vals = [r[2] for r in inside if r[1] == g]
            per_gpu[str(g)] = {"sm_mhz": float(np.median(vals)), "samples": len(vals)}
        return {"sm_mhz": float(np.median(sm)), "sm_max_mhz": float(max(r[3] for r in inside)),
                "reasons": sorted(k for k, bit in NVML_REASONS.items() if mask & bit), "samples": len(sm),
                "source": self.source, "window_ms": 1e3 * (t1 - t0), "per_gpu": per_gpu}


# ---------------------------------------------------------- the B200 model
def build_model(fd, workload, n_t, ic, history="last", world=1, shape=None, periodic=None, dtype=None,
                scaling="weak"):
    """The workload written against the public API (what a user script does).
    With several ranks the grid grows along the slowest axis (weak scaling):
    every rank holds one `shape`-sized slab and passes it as its initial values;
    scaling="strong" keeps the global grid and cuts it into `world` slabs."""
    shape = tuple(shape if shape is not None else WORKLOADS[workload][1])
    if scaling == "weak":
        shape = (shape[0] * world,) + tuple(shape[1:])
    names = ["x", "y", "z"][:len(shape)]
    axes = {"t": np.arange(n_t) * DT}
    for n, a in zip(shape, names):
        axes[a] = np.arange(n) * GRID_SPACING
    periodic = PERIODIC[workload] if periodic is None else periodic
    with contextlib.redirect_stdout(io.StringIO()):
        m = fd.Model(axes, "t", periodic=tuple(names) if periodic else (), history=history,
                     dtype=dtype if dtype is not None else DTYPES.get(workload, np.float64))
        if workload == "c5":
            p = fd.Field(m, "p", n_time_ders=1)
            v = {a: fd.Field(m, "v" + a, edge_axes=(a,), n_time_ders=1) for a in names}
            d_ce = fd.Stencil([-.5, .5], 1, "cell", "edge")
            d_ec = fd.Stencil([-.5, .5], 1, "edge", "cell")
            p.set_IC_values(ic)
            for a in names:
                v[a].set_IC("0")

            def step():
                for a in names:
                    v[a].dot.assign_update((-1 / 1.0) * d_ce.der(p.prev, a))
                    v[a].time_integrate_update()
                div = d_ec.der(v[names[0]].new, names[0])
                for a in names[1:]:
                    div = div + d_ec.der(v[a].new, a)
                p.dot.assign_update((-0.25) * div)
                p.time_integrate_update()
                m.increment_time()
            return m, p, step
        if workload == "c4":
            u = fd.Field(m, "u", n_time_ders=2)
            st = fd.Stencil(list(range(-4, 5)), 2)
            u.set_IC_values(ic)
            u.dot.set_IC("0")

            def step():
                acc = st.der(u.prev, names[0])
                for a in names[1:]:
                    acc = acc + st.der(u.prev, a)
                u.dot.dot.assign_update(0.01 * acc)
                u.dot.time_integrate_update()
                u.time_integrate_update()
                m.increment_time()
            return m, u, step
        T = fd.Field(m, "T", n_time_ders=1)
        st = fd.Stencil([-1, 0, 1], 2)
        T.set_IC_values(ic)
        if not periodic:
            for a in names:
                T.set_BC("0", a, "start")
                T.set_BC("0", a, "end")

    def step():
        acc = st.der(T.prev, names[0])
        for a in names[1:]:
            acc = acc + st.der(T.prev, a)
        T.dot.assign_update(K_DIFF * acc)
        T.time_integrate_update()
        m.increment_time()
    return m, T, step


def fill_ic(view, seed):
    """Uniform [0,1) initial values; levels above 2^27 points repeat one random
    plane along the slowest axis (drawing 10^9 numbers takes the host longer than
    the whole benchmark takes the GPU)."""
    rng = np.random.default_rng(seed)
    if view.size <= (1 << 27):
        view[...] = rng.random(view.shape)
    else:
        view[...] = rng.random(view.shape[1:])[None]


# ------------------------------------------------------------ timed regions
def timed_steps(torch, dist, m, step, K, world, timeline=False):
    """Exactly K steps between two CUDA events on the launching stream.  Every
    queued operation is finished and (several ranks) the ranks are released
    together immediately before the first event: nothing but steps lies between
    the rendezvous and the start of the clock."""
    m.synchronize()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    marks = [torch.cuda.Event(enable_timing=True) for _ in range(K)] if timeline else None
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
        torch.cuda.synchronize()
    t0 = time.monotonic()
    ev0.record()
    for k in range(K):
        step()
        if marks is not None:
            marks[k].record()
    ev1.record()
    t_host = time.monotonic()
    torch.cuda.synchronize()
    t1 = time.monotonic()
    ms = ev0.elapsed_time(ev1)
    out = {"ms": ms, "host_enqueue_ms": 1e3 * (t_host - t0), "window": (t0, t1)}
    if marks is not None:
        out["step_end_ms"] = [ev0.elapsed_time(e) for e in marks]
    return out


def release_garbage(torch, empty_cache=True):
    """A closed model is a cycle of Python objects (model <-> fields <-> recorded
    closures) that still owns device and pinned memory until the cycle collector
    has run; left lying around it made the next workload's host cost per step 2-3x
    higher (profiles/lab/diag_c2_host.py: 0.075 ms fresh, 0.22 ms after a closed
    512^3 model, 0.077 ms after gc.collect())."""
    gc.collect()
    if empty_cache:
        torch.cuda.empty_cache()


def max_over_ranks(torch, dist, value, world):
    if world == 1:
        return float(value)
    t = torch.tensor([value], dtype=torch.float64, device="cuda")
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def run_device_resident(torch, dist, fd, workload, ic, K, W, world, scaling="weak", timeline=False):
    """W warm-up steps, then K timed ones, state resident in HBM."""
    from finite_difference import _device
    m, F, step = build_model(fd, workload, W + K + 2, ic, world=world, scaling=scaling)
    for _ in range(W):
        step()
    launches0 = _device.launch_count()
    got = timed_steps(torch, dist, m, step, K, world, timeline)
    got["launches"] = _device.launch_count() - launches0
    m.synchronize()
    got["stats"] = dict(m._engine.stats)
    got["ms_local"] = got["ms"]
    got["ms"] = max_over_ranks(torch, dist, got["ms"], world)
    if world > 1:
        per_rank = [None] * world
        dist.all_gather_object(per_rank, round(got["ms_local"] / K, 5))
        got["ms_per_step_by_rank"] = per_rank
    got["host_enqueue_ms"] = max_over_ranks(torch, dist, got["host_enqueue_ms"], world)
    m.close()
    del m, F, step
    release_garbage(torch)
    return got


def workload_entry(name, got, K, world, clock, peak, points, scaling="weak"):
    desc, _, bpp = WORKLOADS[name]
    ms_step = got["ms"] / K
    per_gpu_points = points if scaling == "weak" else points / world
    achieved = per_gpu_points * bpp / (ms_step * 1e-3) / 1e9
    total_points = points * world if scaling == "weak" else points
    entry = {"workload": desc, "ms_per_step": ms_step, "point_updates_per_s": total_points / (ms_step * 1e-3),
             "host_enqueue_ms_per_step": got["host_enqueue_ms"] / K,
             "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                          "frac_of_8TBs": achieved / 8000.0, "bytes_per_point_update": bpp,
                          "traffic": recorded_traffic(name)},
             "gpu_launches": got["launches"], "engine_stats": got["stats"]}
    if "ms_per_step_by_rank" in got:
        entry["ms_per_step_by_rank"] = got["ms_per_step_by_rank"]
    if clock is not None:
        entry["clocks"] = got["window"]      # resolved into a summary once the helper has delivered its samples
    return entry


# ------------------------------------------------------- parity self-check
def parity_self_check(fd, world):
    """Small slab-decomposed models against the numpy oracle, bit for bit
    (the oracle is the checker here, nothing of it is timed): heat periodic and
    fixed, wave (4 ghost planes), staggered acoustic fp32.  Slabs are thick
    enough for the overlapped exchange over peer memory."""
    from oracle import fd_numpy as O
    cases = []

    def run(workload, shape, n_t, periodic, ic, dtype=None):
        m, F, step = build_model(fd, workload, n_t, ic, history="all", world=1, shape=shape, periodic=periodic,
                                 dtype=dtype)
        for _ in range(n_t - 1):
            step()
        data = F.data
        stats = dict(m._engine.stats)
        m.close()
        return np.moveaxis(data, 0, 0), stats

    shape3, n_t = (8 * world, 24, 64), 10
    ic = np.random.default_rng(71).random(shape3)
    for periodic in (True, False):
        got, stats = run("c3", shape3, n_t, periodic, ic)
        want, _ = O.run_heat(shape3, n_t, periodic, K_DIFF, DT, ic)
        cases.append(("heat3d_" + ("periodic" if periodic else "fixed"), bool(np.array_equal(got, want, equal_nan=True)),
                      stats.get("peer_halo_exchanges", 0)))
    shape2 = (16 * world, 96)
    ic2 = np.random.default_rng(72).random(shape2)
    got, stats = run("c4", shape2, n_t, True, ic2)
    want, _, _ = O.run_wave(shape2, n_t, 0.01, DT, ic2, np.zeros(shape2))
    cases.append(("wave2d_r4", bool(np.array_equal(got, want, equal_nan=True)), stats.get("peer_halo_exchanges", 0)))
    ic32 = ic.astype(np.float32)
    got, stats = run("c5", shape3, n_t, True, ic32, dtype=np.float32)
    want, _ = O.run_acoustic(shape3, n_t, 1.0, 0.25, DT, ic32, dtype=np.float32)
    cases.append(("acoustic3d_f32", bool(np.array_equal(got, want.astype(np.float64), equal_nan=True)),
                  stats.get("peer_halo_exchanges", 0)))
    return cases


# ------------------------------------------------------------------- arms
def reference_step_fn(workload, shape):
    """(kind, run(n_steps, state) -> (seconds in the step loops, state)) for the
    reference's CPU path: the unmodified FD.py when it is available (staged by
    build() into oracle/_ref, or /root/reference in the build container), else
    the numpy port.  The reference keeps every level of every field, so a long
    run is cut into models of 4 time levels (3 steps) that hand their last level
    on as the next initial condition; only the step loops are timed."""
    from oracle import ref_loader
    periodic = PERIODIC[workload]
    names = ["x", "y", "z"][:len(shape)]
    if workload in ("c2", "c3", "c4") and ref_loader.available():
        ref = ref_loader.load()

        def run(n_steps, state):
            spent, done = 0.0, 0
            while done < n_steps:
                chunk = min(3, n_steps - done)
                axes = {"t": np.arange(chunk + 1) * (1e-2 if workload == "c4" else DT)}
                for n, a in zip(shape, names):
                    axes[a] = np.arange(n) * 1.0
                with contextlib.redirect_stdout(io.StringIO()):
                    m = ref.Model(axes, "t", periodic=tuple(names) if periodic else ())
                    F = ref.Field(m, "F", n_time_ders=2 if workload == "c4" else 1)
                    st = ref.Stencil(list(range(-4, 5)) if workload == "c4" else [-1, 0, 1], 2)
                    F.data[0] = state[0]
                    F.IC_set = True
                    if workload == "c4":
                        F.dot.data[0] = state[1]
                        F.dot.IC_set = True
                    elif not periodic:
                        for k in range(len(names)):
                            sel = [slice(None)] * (len(names) + 1)
                            for idx in (0, -1):
                                sel[1 + k] = idx
                                F.data[tuple(sel)] = 0.0
                    t0 = time.perf_counter()
                    while not m.finished:
                        acc = st.der(F.prev, names[0])
                        for a in names[1:]:
                            acc = acc + st.der(F.prev, a)
                        if workload == "c4":
                            F.dot.dot.assign_update(0.01 * acc)
                            F.dot.time_integrate_update()
                        else:
                            F.dot.assign_update(K_DIFF * acc)
                        F.time_integrate_update()
                        m.increment_time()
                    spent += time.perf_counter() - t0
                state = (np.array(F.data[-1]), np.array(F.dot.data[-1]) if workload == "c4" else None)
                done += chunk
                del m, F
            return spent, state
        return "reference", run

    from oracle import fd_numpy as O

    def run_port(n_steps, state):
        spent, done = 0.0, 0
        while done < n_steps:
            chunk = min(3, n_steps - done)
            t0 = time.perf_counter()
            if workload == "c4":
                u, v, _ = O.run_wave(shape, chunk + 1, 0.01, 1e-2, state[0], state[1])
                state = (np.array(u[-1]), np.array(v[-1]))
            elif workload == "c5":
                p, _ = O.run_acoustic(shape, chunk + 1, 1.0, 0.25, 0.1, state[0])
                state = (np.array(p[-1]), None)        # (velocities restart from zero: timing only)
            else:
                T, _ = O.run_heat(shape, chunk + 1, periodic, K_DIFF, DT, state[0])
                state = (np.array(T[-1]), None)
            spent += time.perf_counter() - t0
            done += chunk
        return spent, state
    return "port", run_port


def cpu_rate(workload, shape, steps_timed=1, warm=1):
    """The reference's CPU path on this host: (kind, point-updates/s, seconds)."""
    kind, run = reference_step_fn(workload, shape)
    ic = np.random.default_rng(0).random(shape)
    state = (ic, np.zeros(shape) if workload == "c4" else None)
    if warm:
        _, state = run(warm, state)
    secs, _ = run(steps_timed, state)
    return kind, float(np.prod(shape)) * steps_timed / secs, secs


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    desc, shape, bpp = WORKLOADS[args.workload]
    if args.workload == "c5":
        shape = (256, 256, 256)          # the reference has no fp32 and 2048^3 fp64 does not fit a host
    # the reference's cost per step does not depend on the step number: at most 24
    # timed steps (about 12 s each at 512^3), so that a K=100 call still ends in minutes
    timed = min(args.steps, 24)
    kind, run = reference_step_fn(args.workload, shape)
    ic = np.random.default_rng(0).random(shape)
    state = (ic, np.zeros(shape) if args.workload == "c4" else None)
    if args.warmup:
        _, state = run(args.warmup, state)
    elapsed, _ = run(timed, state)
    value = float(np.prod(shape)) * timed / elapsed
    sample = (f"{'x'.join(map(str, shape))} grid (the full configuration), {args.warmup} warm-up + {timed} timed steps"
              + (f" (of the {args.steps} requested: per-step cost is constant)" if timed != args.steps else "")
              + ", every level of every field kept in host memory like the reference, models of 4 time levels chained")
    line = {
        "impl": "reference", "metric": "grid-point updates/sec", "value": value, "unit": "point-updates/s",
        "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * elapsed / timed,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": desc, "grid_per_gpu": "x".join(map(str, shape)), "sample": sample, "host_threads": 1,
                   "steps_timed": timed,
                   "code": "unmodified FD.py" if kind == "reference" else "numpy port (oracle/fd_numpy.py)"},
        "cpu_baseline": {"value": value, "unit": "point-updates/s", "cores": 1, "kind": kind, "sample": sample,
                         "host_cpus": os.cpu_count()},
        "e2e": {"value": value, "unit": "point-updates/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line))


def _nccl_options():
    """NCCL kernels on a high-priority stream, so that a halo exchange issued
    while the interior kernel still has blocks to dispatch is not queued behind it."""
    import torch.distributed as dist
    opts = dist.ProcessGroupNCCL.Options()
    opts.is_high_priority_stream = True
    return opts


def run_b200(args):
    import torch
    import torch.distributed as dist
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local), pg_options=_nccl_options())
    import finite_difference as fd
    gc.collect()
    gc.freeze()          # what the imports created stays out of later collections

    desc, shape, bpp = WORKLOADS[args.workload]
    K, W = args.steps, args.warmup
    scaling = args.scaling
    if scaling == "strong" and world > 1:
        shape = (shape[0] // world,) + tuple(shape[1:])     # this rank's share of the fixed global grid
    points = float(np.prod(WORKLOADS[args.workload][1]))
    local_points = float(np.prod(shape))
    peak, peak_src = measured_peak()

    # the helper process samples the clocks of every participating GPU (rank 0 starts it, before any rendezvous)
    clock = None
    if rank == 0:
        uuids = []
        for k in range(world):
            with contextlib.suppress(Exception):
                uuids.append(str(torch.cuda.get_device_properties(k if world > 1 else local).uuid))
        clock = ClockHelper(uuids)

    # ---- several ranks: results first (bit-exact against the oracle), then speed
    parity = None
    if world > 1:
        cases = parity_self_check(fd, world)
        ok = torch.tensor([1 if all(c[1] for c in cases) else 0], device="cuda")
        dist.all_reduce(ok, op=dist.ReduceOp.MIN)
        parity = {"cases": len(cases), "bit_exact": bool(ok.item()), "names": [c[0] for c in cases],
                  "peer_halo_exchanges": [c[2] for c in cases]}
        if not parity["bit_exact"]:
            sys.stderr.write(f"rank {rank}: parity self-check FAILED: {cases}\n")
            dist.destroy_process_group()
            sys.exit(3)

    # ---- pinned host buffer for initial values: one allocation, viewed per workload
    todo = [args.workload]
    if not args.no_workloads:
        todo += [w for w in (("c2", "c4", "c5") if world == 1 else ("c5",)) if w != args.workload]
    need = max(int(np.prod(WORKLOADS[w][1] if not (w == args.workload and scaling == "strong") else shape))
               * np.dtype(DTYPES.get(w, np.float64)).itemsize for w in todo)
    pinned_bytes = torch.empty(need, dtype=torch.uint8, pin_memory=True)

    def host_view(w, shp):
        dt = np.dtype(DTYPES.get(w, np.float64))
        n = int(np.prod(shp))
        return pinned_bytes[:n * dt.itemsize].numpy().view(dt).reshape(shp)

    # ---- headline: device-resident run -------------------------------------
    ic = host_view(args.workload, shape)
    fill_ic(ic, rank)
    head = run_device_resident(torch, dist, fd, args.workload, ic, K, W, world, scaling=scaling,
                               timeline=bool(args.timeline))
    ms = head["ms"]
    if args.timeline:
        rows = [None] * world
        mine = {"rank": rank, "ms": head["ms_local"], "step_end_ms": head["step_end_ms"]}
        if world > 1:
            dist.all_gather_object(rows, mine)
        else:
            rows = [mine]
        if rank == 0:
            with open(args.timeline, "w") as f:
                json.dump({"n_gpus": world, "steps": K, "warmup": W, "workload": args.workload, "scaling": scaling,
                           "ranks": rows}, f)

    # ---- end to end through the public API with host buffers -------------
    np_dtype = DTYPES.get(args.workload, np.float64)
    result_pinned = torch.empty(int(local_points), dtype=torch.float32 if np_dtype == np.float32 else torch.float64,
                                pin_memory=True)   # (empty_like would not be pinned)
    result = result_pinned.numpy().reshape(shape)

    def end_to_end(steps, history="last"):
        t0 = time.perf_counter()
        m2, F2, step2 = build_model(fd, args.workload, steps + 1, ic, history=history, world=world, scaling=scaling)
        for _ in range(steps):
            step2()
        out = F2.local_level(steps, out=result)   # device -> pinned host buffer, synchronous
        m2.synchronize()
        torch.cuda.synchronize()
        secs = time.perf_counter() - t0
        m2.close()
        del m2, F2, step2
        release_garbage(torch, empty_cache=False)     # (the allocator's pools stay warm for the timed run)
        return secs, out

    end_to_end(W)                            # warm-up: allocator pools (device ring, pinned host level)
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
        torch.cuda.synchronize()
    e2e_s, final = end_to_end(K)
    e2e_s = max_over_ranks(torch, dist, e2e_s, world)
    assert final.shape == tuple(shape) and np.isfinite(final[::7]).all()
    bytes_level = local_points * np.dtype(np_dtype).itemsize

    # ---- the reference-default history policy, end to end (one GPU) --------
    history_legs = None
    if world == 1 and args.workload == "c3" and not args.no_history_legs:
        history_legs = {}
        try:
            import psutil
            free_gib = psutil.virtual_memory().available / 2 ** 30
        except Exception:
            free_gib = 0.0
        hk = 8
        for label, policy in (("c3_history_last", "last"), ("c3_history_every_8", 8), ("c3_history_all", "all")):
            need_gib = {"last": 0, 8: 2 * 2, "all": 2 * hk}[policy] * bytes_level / 2 ** 30 + 8
            if free_gib < need_gib:
                history_legs[label] = {"skipped": f"needs {need_gib:.0f} GiB of free host memory, {free_gib:.0f} available"}
                continue
            end_to_end(2, policy)
            secs, _ = end_to_end(hk, policy)
            history_legs[label] = {"steps": hk, "seconds": secs, "ms_per_step": 1e3 * secs / hk,
                                   "point_updates_per_s": local_points * hk / secs,
                                   "d2h_bytes_per_step": {"last": bytes_level / hk, 8: 2 * 2 * bytes_level / hk,
                                                          "all": 2 * bytes_level}[policy],
                                   "policy": f"Model(history={policy!r}); T and T.dot levels kept"}
        release_garbage(torch)

    # ---- the other configurations, same process, same timing rules ---------
    others = {}
    for w in todo[1:]:
        wshape = WORKLOADS[w][1]
        wic = host_view(w, wshape)
        fill_ic(wic, rank)
        got = run_device_resident(torch, dist, fd, w, wic, K, W, world)
        others[w] = workload_entry(w, got, K, world, clock, peak, float(np.prod(wshape)))
        if world > 1:
            others[w]["grid"] = f"{wshape[0] * world}x{wshape[1]}x{wshape[2]} over {world} GPUs"
    # one slab alone on rank 0's GPU right after (same box, same process): the
    # denominator of the weak-scaling efficiency of this very run
    alone = {}
    if world > 1 and not args.no_workloads:
        for w in dict.fromkeys([args.workload] + list(others)):
            if w == args.workload and scaling == "strong":
                continue
            if rank == 0:
                os.environ["FD_B200_DIST"] = "0"
                try:
                    wic = host_view(w, WORKLOADS[w][1])
                    got = run_device_resident(torch, None, fd, w, wic, K, W, 1)
                    alone[w] = got["ms"] / K
                finally:
                    del os.environ["FD_B200_DIST"]
            dist.barrier()

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return

    total_points = points * world if scaling == "weak" else points
    value = total_points * K / (ms * 1e-3)
    achieved = local_points * bpp / (ms / K * 1e-3) / 1e9
    stats = head["stats"]
    line = {
        "metric": "grid-point updates/sec", "value": value, "unit": "point-updates/s", "n_gpus": world,
        "steps": K, "warmup": W, "ms_per_step": ms / K, "higher_is_better": True, "scaling": scaling,
        "vs_baseline": None, "dtype": "f32" if np_dtype == np.float32 else "f64", "data": "synthetic",
        "config": {"workload": desc, "grid_per_gpu": "x".join(map(str, shape)),
                   "l2": "inputs larger than L2 (state level = %.0f MiB vs 126 MB L2)" % (bytes_level / 2 ** 20),
                   "multi_gpu": (f"slab decomposition along x over {world} ranks, halo planes "
                                 + ("written into the neighbours' memory over NVLink (peer-memory push + flag words)"
                                    if stats.get("peer_halo_exchanges") else "exchanged with NCCL send/recv")
                                 + " every step" if world > 1 else "single GPU"),
                   "history": "last", "engine_stats": stats},
        "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                     "traffic": recorded_traffic(args.workload), "peak_source": peak_src,
                     "frac_of_8TBs": achieved / 8000.0, "bytes_per_point_update": bpp},
        "e2e": {"value": total_points * K / e2e_s, "unit": "point-updates/s",
                "h2d_bytes_per_step": bytes_level / K, "d2h_bytes_per_step": bytes_level / K,
                "seconds": e2e_s,
                "includes": "model construction, IC upload from pinned host memory, K steps, final level read back into a pinned host buffer"},
        "gpu_launches": head["launches"],
        "host_enqueue_ms_per_step": head["host_enqueue_ms"] / K,
        "clocks": head["window"],
    }
    if "ms_per_step_by_rank" in head:
        line["ms_per_step_by_rank"] = head["ms_per_step_by_rank"]
    if parity is not None:
        line["parity_check"] = parity
    if others:
        line["workloads"] = others
    if history_legs:
        line.setdefault("workloads", {}).update(history_legs)
    if alone:
        line["single_gpu_same_run"] = {w: {"ms_per_step": t,
                                           "weak_efficiency": t / (ms / K if w == args.workload else others[w]["ms_per_step"])}
                                       for w, t in alone.items()}
    if world == 1 and not args.no_cpu_baseline:
        sshape = WORKLOADS[args.workload][1] if args.workload != "c5" else (256, 256, 256)
        kind, rate, secs = cpu_rate(args.workload, sshape)
        line["cpu_baseline"] = {"value": rate, "unit": "point-updates/s", "cores": 1, "kind": kind,
                                "sample": f"{'x'.join(map(str, sshape))} grid, 1 warm-up + 1 timed step "
                                          f"({secs:.1f} s), full numpy history like the reference",
                                "host_cpus": os.cpu_count()}
    line["clocks"] = clock.summary(line["clocks"])
    for entry in line.get("workloads", {}).values():
        if isinstance(entry.get("clocks"), tuple):
            entry["clocks"] = clock.summary(entry["clocks"])
    print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


def main():
    if len(sys.argv) >= 3 and sys.argv[1] == "--clock-helper":
        clock_helper_main([u for u in sys.argv[2].split(",") if u])
        return
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=100)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default="c3", choices=sorted(WORKLOADS))
    ap.add_argument("--scaling", default="weak", choices=["weak", "strong"],
                    help="several GPUs: one full grid per GPU (weak, default) or the one grid cut into slabs (strong)")
    ap.add_argument("--no-workloads", action="store_true", help="headline workload only")
    ap.add_argument("--no-history-legs", action="store_true")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--timeline", default=None, help="write the end time of every step on every rank to this JSON file")
    ap.add_argument("--dx", type=float, default=1.0, help="grid spacing of every spatial axis (default 1, the benchmark definition)")
    args = ap.parse_args()
    global GRID_SPACING
    GRID_SPACING = args.dx
    args.warmup = max(args.warmup, 3) if args.impl == "b200" else args.warmup
    if args.impl == "reference":
        run_reference(args)
    else:
        run_b200(args)


if __name__ == "__main__":
    main()
