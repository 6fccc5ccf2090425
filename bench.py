#!/usr/bin/env python
"""Benchmark of the finite_difference time-stepping hot path on B200.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference] [--workload c3|c2|c4]

One "step" is one time step of the model (every grid point advanced once).
Default workload (N=1): BASELINE.json configs[2], 3-D heat diffusion 512^3
fp64, 7-point stencil, periodic -- the configuration the north-star target is
quoted on.  Prints ONE JSON line (see DESIGN.md "Measurement").

  value      point-updates/s, state resident in HBM, CUDA events on the
             launching stream around exactly K steps driven through the public
             `finite_difference` API (the recorded-statement layer included)
  e2e        the same K steps through the public API with HOST buffers:
             initial condition uploaded from pinned host memory, final level
             read back to the host, both inside the timed region
  roofline   algorithmic bytes (16 B/point: one read + one write of the state)
             over the measured step time, against the measured HBM copy peak
  cpu_baseline  the numpy port of the reference path (oracle/fd_numpy.py) on
             a bounded sample, on this box's host cores (numpy: 1 core)
"""

from __future__ import annotations

import argparse
import contextlib
import io
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
for p in (ROOT, os.path.join(ROOT, "finite-difference_b200")):
    if p not in sys.path:
        sys.path.insert(0, p)

WORKLOADS = {
    # name: (description, spatial shape, bytes per point update)
    "c3": ("3-D heat diffusion 512^3 fp64, 7-point stencil, periodic (BASELINE.json configs[2])", (512, 512, 512), 16),
    "c2": ("2-D heat diffusion 8192^2 fp64, 5-point stencil, fixed BCs (BASELINE.json configs[1])", (8192, 8192), 16),
    "c4": ("2-D wave equation 8192^2 fp64, n_time_ders=2, 9-point-per-axis stencils, periodic (BASELINE.json configs[3])",
           (8192, 8192), 32),
    "c5": ("3-D staggered cell/edge acoustic model fp32, periodic, 256x2048x2048 slab per GPU "
           "(2048^3 on 8 GPUs, BASELINE.json configs[4])", (256, 2048, 2048), 32),
}
DTYPES = {"c5": np.float32}
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


class ClockSampler:
    """SM clock and throttle reasons during the timed region.

    NVML in-process (a thread polling every 2 ms: the timed region of the default
    run is only ~40 ms long, shorter than the start-up of an nvidia-smi process);
    `nvidia-smi -lms` as the fallback when the NVML binding is unavailable."""
    QUERY = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
             "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")
    NAMES = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
    NVML_BITS = {"hw_slowdown": 0x8, "hw_thermal_slowdown": 0x40, "sw_thermal_slowdown": 0x20, "sw_power_cap": 0x4}

    def __init__(self, index=0):
        self.index, self.rows, self.proc, self.thread = index, [], None, None
        self.nvml, self.handle, self.stop, self.source = None, None, threading.Event(), None

    def _nvml_handle(self):
        import pynvml
        pynvml.nvmlInit()
        handle = None
        try:
            import torch
            uuid = str(torch.cuda.get_device_properties(self.index).uuid)
            handle = pynvml.nvmlDeviceGetHandleByUUID(uuid if uuid.startswith("GPU-") else "GPU-" + uuid)
        except Exception:
            handle = pynvml.nvmlDeviceGetHandleByIndex(self.index)
        pynvml.nvmlDeviceGetClockInfo(handle, pynvml.NVML_CLOCK_SM)       # fails here rather than in the thread
        return pynvml, handle

    def __enter__(self):
        try:
            self.nvml, self.handle = self._nvml_handle()
            self.source = "nvml"
            self.thread = threading.Thread(target=self._poll_nvml, daemon=True)
            self.thread.start()
            return self
        except Exception:
            self.nvml = None
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", f"--query-gpu={self.QUERY}", "--format=csv,noheader,nounits", "-lms", "100",
                 "-i", str(self.index)], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.source = "nvidia-smi"
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
        except OSError:
            self.proc = None
        return self

    def _poll_nvml(self):
        n = self.nvml
        try:
            mx = n.nvmlDeviceGetMaxClockInfo(self.handle, n.NVML_CLOCK_SM)
        except Exception:
            mx = None
        while not self.stop.is_set():
            try:
                sm = n.nvmlDeviceGetClockInfo(self.handle, n.NVML_CLOCK_SM)
                try:
                    mask = n.nvmlDeviceGetCurrentClocksEventReasons(self.handle)
                except Exception:
                    mask = n.nvmlDeviceGetCurrentClocksThrottleReasons(self.handle)
                self.rows.append([str(sm), str(mx if mx is not None else sm)]
                                 + ["Active" if mask & self.NVML_BITS[k] else "Not Active" for k in self.NAMES])
            except Exception:
                pass
            self.stop.wait(0.002)

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def __exit__(self, *exc):
        if self.nvml is not None:
            self.stop.set()
            self.thread.join(timeout=2)
        elif self.proc is not None:
            time.sleep(0.15)
            self.proc.terminate()
            self.thread.join(timeout=2)

    def summary(self):
        sm, mx, reasons = [], [], set()
        for r in self.rows:
            try:
                sm.append(float(r[0]))
                mx.append(float(r[1]))
            except (ValueError, IndexError):
                continue
            for name, v in zip(self.NAMES, r[2:6]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        if not sm:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0, "source": self.source}
        return {"sm_mhz": float(np.median(sm)), "sm_max_mhz": float(max(mx)), "reasons": sorted(reasons),
                "samples": len(sm), "source": self.source}


# ---------------------------------------------------------- the B200 model
def build_model(fd, workload, n_t, ic, history="last", world=1):
    """The workload written against the public API (what a user script does).
    With several ranks the grid grows along the slowest axis (weak scaling):
    every rank holds one `shape`-sized slab and passes it as its initial values."""
    shape = WORKLOADS[workload][1]
    shape = (shape[0] * world,) + tuple(shape[1:])
    names = ["x", "y", "z"][:len(shape)]
    axes = {"t": np.arange(n_t) * DT}
    for n, a in zip(shape, names):
        axes[a] = np.arange(n) * GRID_SPACING
    periodic = workload in ("c3", "c4", "c5")
    with contextlib.redirect_stdout(io.StringIO()):
        m = fd.Model(axes, "t", periodic=tuple(names) if periodic else (), history=history,
                     dtype=DTYPES.get(workload, np.float64))
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
                div = d_ec.der(v["x"].new, "x") + d_ec.der(v["y"].new, "y") + d_ec.der(v["z"].new, "z")
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


def cpu_port_rate(workload, shape, steps_timed=1):
    """Numpy port of the reference path on this host: point-updates/s."""
    from oracle import fd_numpy as O
    ic = np.random.default_rng(0).random(shape)
    # a 1-step run first as warm-up (page faults, numpy set-up), then the timed run
    if workload == "c4":
        O.run_wave(shape, 2, 0.01, 1e-2, ic, np.zeros(shape))
        t0 = time.perf_counter()
        O.run_wave(shape, 1 + steps_timed, 0.01, 1e-2, ic, np.zeros(shape))
    else:
        periodic = workload == "c3"
        O.run_heat(shape, 2, periodic, K_DIFF, DT, ic)
        t0 = time.perf_counter()
        O.run_heat(shape, 1 + steps_timed, periodic, K_DIFF, DT, ic)
    dt = time.perf_counter() - t0
    return float(np.prod(shape)) * steps_timed / dt, dt


def sample_shape(workload, target_points):
    full = WORKLOADS[workload][1]
    if np.prod(full) <= target_points:
        return full
    side = int(round(target_points ** (1.0 / len(full))))
    side = max(32, side - side % 32)
    return tuple(min(side, n) for n in full)



def _nccl_options():
    """NCCL kernels on a high-priority stream, so that a halo exchange issued
    while the interior kernel still has blocks to dispatch is not queued behind it."""
    import torch.distributed as dist
    opts = dist.ProcessGroupNCCL.Options()
    opts.is_high_priority_stream = True
    return opts


# ------------------------------------------------------------------- arms
def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    desc, shape, bpp = WORKLOADS[args.workload]
    # bounded sample: K+W steps of the numpy port must end within a few minutes
    rate_guess = 1.0e7
    budget_s = 150.0
    pts = rate_guess * budget_s / max(1, args.steps + args.warmup)
    sshape = sample_shape(args.workload, pts)
    from oracle import fd_numpy as O
    ic = np.random.default_rng(0).random(sshape)
    periodic = args.workload in ("c3", "c4")

    def run(n_steps):
        if args.workload == "c4":
            O.run_wave(sshape, n_steps + 1, 0.01, 1e-2, ic, np.zeros(sshape))
        else:
            O.run_heat(sshape, n_steps + 1, periodic, K_DIFF, DT, ic)
    run(args.warmup)
    t0 = time.perf_counter()
    run(args.steps)
    elapsed = time.perf_counter() - t0
    value = float(np.prod(sshape)) * args.steps / elapsed
    sample = f"{'x'.join(map(str, sshape))} grid (same stencil, dtype and boundary kind), {args.steps} steps, full numpy history"
    line = {
        "impl": "reference", "metric": "grid-point updates/sec", "value": value, "unit": "point-updates/s",
        "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * elapsed / args.steps,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": desc, "sample": sample, "host_threads": 1},
        "cpu_baseline": {"value": value, "unit": "point-updates/s", "cores": 1, "kind": "port", "sample": sample,
                         "host_cpus": os.cpu_count()},
        "e2e": {"value": value, "unit": "point-updates/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line))


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
    from finite_difference import _device

    desc, shape, bpp = WORKLOADS[args.workload]
    points = float(np.prod(shape))
    K, W = args.steps, args.warmup

    # ---- device-resident run --------------------------------------------
    np_dtype = DTYPES.get(args.workload, np.float64)
    pinned = torch.empty(int(points), dtype=torch.float32 if np_dtype == np.float32 else torch.float64,
                         pin_memory=True)
    ic = pinned.numpy().reshape(shape)
    ic[...] = np.random.default_rng(rank).random(shape)
    m, F, step = build_model(fd, args.workload, W + K + 2, ic, world=world)
    for _ in range(W):
        step()
    m.synchronize()
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    launches0 = _device.launch_count()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    with ClockSampler(local) as clocks:
        ev0.record()
        for _ in range(K):
            step()
        ev1.record()
        torch.cuda.synchronize()
    ms = ev0.elapsed_time(ev1)
    launches = _device.launch_count() - launches0
    m.synchronize()
    stats = dict(m._engine.stats)
    if world > 1:
        t = torch.tensor([ms], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t.item())
    del m, F, step
    torch.cuda.empty_cache()

    # ---- end to end through the public API with host buffers -------------
    result_pinned = torch.empty(pinned.shape, dtype=pinned.dtype, pin_memory=True)   # (empty_like would not be pinned)
    result = result_pinned.numpy().reshape(shape)

    def end_to_end(steps):
        t0 = time.perf_counter()
        m2, F2, step2 = build_model(fd, args.workload, steps + 1, ic, world=world)
        for _ in range(steps):
            step2()
        out = F2.local_level(steps, out=result)   # device -> pinned host buffer, synchronous
        m2.synchronize()
        torch.cuda.synchronize()
        return time.perf_counter() - t0, out

    end_to_end(W)                            # warm-up: allocator pools (device ring, pinned host level)
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    e2e_s, final = end_to_end(K)
    if world > 1:
        t = torch.tensor([e2e_s], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        e2e_s = float(t.item())
    assert final.shape == tuple(shape) and np.isfinite(final[::7]).all()
    bytes_level = points * np.dtype(np_dtype).itemsize

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return

    value = world * points * K / (ms * 1e-3)
    peak, peak_src = measured_peak()
    achieved = points * bpp / (ms / K * 1e-3) / 1e9
    line = {
        "metric": "grid-point updates/sec", "value": value, "unit": "point-updates/s", "n_gpus": world,
        "steps": K, "warmup": W, "ms_per_step": ms / K, "higher_is_better": True, "scaling": "weak",
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
        "e2e": {"value": world * points * K / e2e_s, "unit": "point-updates/s",
                "h2d_bytes_per_step": bytes_level / K, "d2h_bytes_per_step": bytes_level / K,
                "seconds": e2e_s,
                "includes": "model construction, IC upload from pinned host memory, K steps, final level read back into a pinned host buffer"},
        "gpu_launches": launches,
        "clocks": clocks.summary(),
    }
    if world == 1 and not args.no_cpu_baseline:
        sshape = sample_shape(args.workload, 1.4e8 if args.workload == "c3" else 6.8e7)
        rate, secs = cpu_port_rate(args.workload, sshape)
        line["cpu_baseline"] = {"value": rate, "unit": "point-updates/s", "cores": 1, "kind": "port",
                                "sample": f"{'x'.join(map(str, sshape))} grid, 1 warm-up + 1 timed step "
                                          f"({secs:.1f} s), full numpy history like the reference",
                                "host_cpus": os.cpu_count()}
    print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=100)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default="c3", choices=sorted(WORKLOADS))
    ap.add_argument("--no-cpu-baseline", action="store_true")
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
