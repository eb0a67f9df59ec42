#!/usr/bin/env python3
"""Host <-> device copy ceiling of this box with N ranks copying AT THE SAME TIME (the limit of bench.py's
`e2e` at N > 1: every rank streams its x / top_error in and y / dL/dx out through one host).

  python tools/pcie_probe.py                                   one rank
  python -m torch.distributed.run --nproc-per-node N --master-addr 127.0.0.1 tools/pcie_probe.py [--bind 1]

Per rank: pinned host buffers of --mb MiB, one copy stream per direction; phases H2D only, D2H only, both
directions at once; all ranks start each phase together (barrier) and the phase time is the max over
ranks.  --bind 1 pins every rank (and the pages it first touches) to its own slice of the host cores;
--register 1 pins page-aligned buffers with cudaHostRegister instead of torch's pin_memory allocator.
Rank 0 prints one JSON line: GB/s per rank and aggregate for each phase, plus the host's CPU / NUMA layout.
"""
import argparse
import json
import os
import subprocess
import time

import torch
import torch.distributed as dist


def host_layout():
    out = {"cpus": os.cpu_count(), "affinity": len(os.sched_getaffinity(0))}
    try:
        txt = subprocess.run(["lscpu"], capture_output=True, text=True, timeout=5).stdout
        for line in txt.splitlines():
            k = line.split(":")[0].strip()
            if k in ("Model name", "Socket(s)", "NUMA node(s)", "Thread(s) per core", "Core(s) per socket"):
                out[k] = line.split(":", 1)[1].strip()
    except Exception:
        pass
    try:
        for line in open("/proc/meminfo"):
            if line.startswith("MemTotal"):
                out["mem_total_gb"] = round(int(line.split()[1]) / 1e6, 1)
    except Exception:
        pass
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--mb", type=int, default=512)
    ap.add_argument("--reps", type=int, default=6)
    ap.add_argument("--bind", type=int, default=0)
    ap.add_argument("--register", type=int, default=0)
    a = ap.parse_args()
    rank = int(os.environ.get("RANK", "0")); world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if a.bind:
        cores = sorted(os.sched_getaffinity(0))
        per = max(1, len(cores) // world)
        os.sched_setaffinity(0, set(cores[local * per:(local + 1) * per]))
    dev = torch.device("cuda", local)
    torch.cuda.set_device(dev)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)
    n = a.mb * 1024 * 1024 // 4

    def host_buf():
        if a.register:
            t = torch.empty(n + 1024, dtype=torch.float32)
            off = (-t.data_ptr() // 4) % 1024                     # 4 KiB alignment
            t = t[off:off + n]
            t.fill_(1.0)                                         # first touch on this rank's cores
            rc = torch.cuda.cudart().cudaHostRegister(t.data_ptr(), n * 4, 0)
            assert int(rc) == 0, rc
            return t
        t = torch.empty(n, dtype=torch.float32).pin_memory()
        t.fill_(1.0)
        return t

    hin, hout = host_buf(), host_buf()
    din, dout = torch.empty(n, device=dev), torch.ones(n, device=dev)
    s_in, s_out = torch.cuda.Stream(dev), torch.cuda.Stream(dev)

    def phase(do_in, do_out):
        def once():
            if do_in:
                with torch.cuda.stream(s_in):
                    din.copy_(hin, non_blocking=True)
            if do_out:
                with torch.cuda.stream(s_out):
                    hout.copy_(dout, non_blocking=True)
        once()
        torch.cuda.synchronize(dev)
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize(dev)
        t0 = time.perf_counter()
        for _ in range(a.reps):
            once()
        torch.cuda.synchronize(dev)
        dt = torch.tensor([(time.perf_counter() - t0) / a.reps], device=dev)
        if world > 1:
            dist.all_reduce(dt, op=dist.ReduceOp.MAX)
        return n * 4 / float(dt.item()) / 1e9                     # GB/s per rank and direction (slowest rank)

    res = {}
    for name, di, do in (("h2d_only", True, False), ("d2h_only", False, True), ("both", True, True)):
        g = phase(di, do)
        res[name] = {"gbs_per_rank_per_direction": round(g, 2),
                     "gbs_aggregate": round(g * world * (2 if (di and do) else 1), 1)}
    if rank == 0:
        print(json.dumps({"tool": "pcie_probe", "ranks": world, "mb_per_buffer": a.mb, "bind_cores": bool(a.bind),
                          "cudaHostRegister": bool(a.register), "phases": res, "host": host_layout()}))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
