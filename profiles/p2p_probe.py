"""Two (or more) ranks: halo exchange over peer memory through the C ABI
(fd_ipc_* / fd_halo_push / fd_halo_wait) against NCCL isend/irecv.

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 \
        --master-port 29533 profiles/p2p_probe.py
"""
import ctypes as C
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "finite-difference_b200")]
import torch
import torch.distributed as dist

from finite_difference import _native as nat


def main():
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", rank)))
    dist.init_process_group("nccl")
    lib = nat.lib
    plane = 512 * 512                                  # one fp64 plane of the 512^3 benchmark
    owned = 8
    level = torch.zeros((owned + 2) * plane, dtype=torch.float64, device="cuda")   # [ghost | owned | ghost]
    flags = torch.zeros(8, dtype=torch.int64, device="cuda")
    err = torch.zeros(4, dtype=torch.int64, device="cuda")
    torch.cuda.synchronize()

    def export(t):
        h = C.create_string_buffer(nat.FD_IPC_HANDLE_BYTES)
        off = C.c_longlong()
        nat.check(lib.fd_ipc_export(t.data_ptr(), h, C.byref(off)))
        return bytes(h.raw), off.value

    mine = {"level": export(level), "flags": export(flags)}
    everyone = [None] * world
    dist.all_gather_object(everyone, mine)
    prev, nxt = (rank - 1) % world, (rank + 1) % world
    opened = {}

    def peer_ptr(r, key):
        handle, off = everyone[r][key]
        if (r, handle) not in opened:
            base = C.c_void_p()
            nat.check(lib.fd_ipc_open(handle, C.byref(base)))
            opened[(r, handle)] = base.value
        return opened[(r, handle)] + off

    esz = 8
    # my low owned plane -> prev's high ghost; my high owned plane -> next's low ghost
    src_lo = level.data_ptr() + 1 * plane * esz
    src_hi = level.data_ptr() + owned * plane * esz
    dst_lo = peer_ptr(prev, "level") + (owned + 1) * plane * esz
    dst_hi = peer_ptr(nxt, "level") + 0
    # flag words in the receiver: [0] = "your low ghost is fresh", [1] = "your high ghost is fresh"
    flag_lo = peer_ptr(prev, "flags") + 1 * 8
    flag_hi = peer_ptr(nxt, "flags") + 0 * 8
    my_flag_lo, my_flag_hi = flags.data_ptr(), flags.data_ptr() + 8
    stream = torch.cuda.current_stream().cuda_stream
    dist.barrier()

    ok = True
    for step in range(1, 4):
        level[plane:(owned + 1) * plane] = float(rank * 100 + step)
        nat.check(lib.fd_halo_push(src_lo, dst_lo, src_hi, dst_hi, plane * esz, flag_lo, flag_hi, step, stream))
        nat.check(lib.fd_halo_wait(my_flag_lo, my_flag_hi, step, 5_000_000_000, err.data_ptr(), stream))
        torch.cuda.synchronize()
        lo_ghost, hi_ghost = level[:plane], level[(owned + 1) * plane:]
        want_lo, want_hi = float(prev * 100 + step), float(nxt * 100 + step)
        good = bool((lo_ghost == want_lo).all() and (hi_ghost == want_hi).all() and int(err.sum()) == 0)
        ok &= good
        dist.barrier()
    print(f"rank {rank}: peer-memory halo exchange {'correct' if ok else 'WRONG'}", flush=True)

    def timed(fn, n=300):
        for _ in range(20):
            fn()
        torch.cuda.synchronize()
        dist.barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(n):
            fn()
        e1.record()
        torch.cuda.synchronize()
        return e0.elapsed_time(e1) / n * 1e3

    counter = [100]

    def p2p():
        counter[0] += 1
        lib.fd_halo_push(src_lo, dst_lo, src_hi, dst_hi, plane * esz, flag_lo, flag_hi, counter[0], stream)
        lib.fd_halo_wait(my_flag_lo, my_flag_hi, counter[0], 5_000_000_000, err.data_ptr(), stream)

    lo_own, hi_own = level[plane:2 * plane], level[owned * plane:(owned + 1) * plane]
    lo_ghost, hi_ghost = level[:plane], level[(owned + 1) * plane:]

    def nccl():
        ops = [dist.P2POp(dist.isend, lo_own, prev), dist.P2POp(dist.isend, hi_own, nxt),
               dist.P2POp(dist.irecv, hi_ghost, nxt), dist.P2POp(dist.irecv, lo_ghost, prev)]
        for r in dist.batch_isend_irecv(ops):
            r.wait()

    t_p2p = timed(p2p)
    t_nccl = timed(nccl)
    t0 = time.perf_counter()
    for _ in range(300):
        p2p()
    host_p2p = (time.perf_counter() - t0) / 300 * 1e6
    torch.cuda.synchronize()
    dist.barrier()
    t0 = time.perf_counter()
    for _ in range(300):
        nccl()
    host_nccl = (time.perf_counter() - t0) / 300 * 1e6
    torch.cuda.synchronize()
    if rank == 0:
        print(f"2 MiB planes both ways, world {world}: peer-memory push+wait {t_p2p:.1f} us (host {host_p2p:.1f} us), "
              f"NCCL batch_isend_irecv {t_nccl:.1f} us (host {host_nccl:.1f} us); timeouts {int(err.sum())}", flush=True)
    dist.barrier()
    dist.destroy_process_group()


main()
