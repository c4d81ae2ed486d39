#!/usr/bin/env python
"""What bounds the end-to-end (host-buffer) leg when several GPUs copy at once.

Every rank copies between its GPU and page-locked host memory at the same time as all the others (barrier before each
timed section) and reports GB/s for H2D alone, D2H alone and both directions together, for three kinds of host
buffer -- plain cudaHostAlloc, write-combined cudaHostAlloc (H2D source only), and 2 MB-aligned malloc + MADV_HUGEPAGE
+ cudaHostRegister -- and three copy sizes.  Rank 0 prints one JSON object (min / mean over ranks) plus what the box
says about itself (cores, NUMA nodes, PCIe link, GPU topology).

    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port 29533 tools/e2e_probe.py
"""
import ctypes as C
import json
import mmap
import os
import subprocess
import sys

import torch
import torch.distributed as dist


def sh(cmd):
    try:
        return subprocess.run(cmd, shell=True, capture_output=True, text=True, timeout=20).stdout.strip()
    except Exception as e:  # noqa: BLE001
        return "n/a (%s)" % e


def main():
    rank = int(os.environ.get("RANK", 0))
    world = int(os.environ.get("WORLD_SIZE", 1))
    local = int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    dev = torch.device("cuda", local)
    rt = C.CDLL("libcudart.so.12")
    rt.cudaHostAlloc.argtypes = [C.POINTER(C.c_void_p), C.c_size_t, C.c_uint]
    rt.cudaHostRegister.argtypes = [C.c_void_p, C.c_size_t, C.c_uint]
    rt.cudaMemcpyAsync.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t, C.c_int, C.c_void_p]
    NB = 1 << 30
    d_a = torch.empty(NB, dtype=torch.uint8, device=dev)
    d_b = torch.empty(NB, dtype=torch.uint8, device=dev)
    s_up, s_down = torch.cuda.Stream(), torch.cuda.Stream()

    def host_plain():
        p = C.c_void_p()
        assert rt.cudaHostAlloc(C.byref(p), NB, 0) == 0
        return p.value

    def host_wc():
        p = C.c_void_p()
        assert rt.cudaHostAlloc(C.byref(p), NB, 4) == 0                # cudaHostAllocWriteCombined
        return p.value

    def host_huge():
        m = mmap.mmap(-1, NB + (2 << 20), flags=mmap.MAP_PRIVATE | mmap.MAP_ANONYMOUS)
        base = C.addressof(C.c_char.from_buffer(m))
        p = (base + (2 << 20) - 1) & ~((2 << 20) - 1)
        libc = C.CDLL("libc.so.6")
        libc.madvise.argtypes = [C.c_void_p, C.c_size_t, C.c_int]
        libc.madvise(C.c_void_p(p), NB, 14)                             # MADV_HUGEPAGE
        C.memset(C.c_void_p(p), 1, NB)                                  # fault the pages in (as huge pages where the kernel allows)
        rc = rt.cudaHostRegister(C.c_void_p(p), NB, 0)
        host_huge.keep = m
        return p if rc == 0 else None

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()

    def timed(fn, reps=3):
        best = 0.0
        for _ in range(reps):
            barrier()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            n = fn()
            s_up.synchronize()
            s_down.synchronize()
            e1.record()
            torch.cuda.synchronize()
            best = max(best, n / e0.elapsed_time(e1) / 1e6)
        return best

    def copies(h_up, h_down, chunk, up=True, down=True):
        def fn():
            n = 0
            for off in range(0, NB, chunk):
                if up:
                    rt.cudaMemcpyAsync(C.c_void_p(d_a.data_ptr() + off), C.c_void_p(h_up + off), chunk, 1, C.c_void_p(s_up.cuda_stream))
                    n += chunk
                if down:
                    rt.cudaMemcpyAsync(C.c_void_p(h_down + off), C.c_void_p(d_b.data_ptr() + off), chunk, 2, C.c_void_p(s_down.cuda_stream))
                    n += chunk
            return n
        return fn

    plain_a, plain_b = host_plain(), host_plain()
    C.memset(C.c_void_p(plain_a), 1, NB)
    C.memset(C.c_void_p(plain_b), 1, NB)
    wc = host_wc()
    huge = host_huge()
    res = {}
    for chunk_mb in (16, 64, 256):
        chunk = chunk_mb << 20
        row = {"h2d": timed(copies(plain_a, plain_b, chunk, True, False)), "d2h": timed(copies(plain_a, plain_b, chunk, False, True)),
               "both_sum": timed(copies(plain_a, plain_b, chunk, True, True)),
               "h2d_write_combined": timed(copies(wc, plain_b, chunk, True, False))}
        if huge:
            row["h2d_hostregister_2mb"] = timed(copies(huge, plain_b, chunk, True, False))
            row["d2h_hostregister_2mb"] = timed(copies(plain_a, huge, chunk, False, True))
        res["chunk_%dMB" % chunk_mb] = row
    keys = sorted({k for r in res.values() for k in r})
    flat = torch.tensor([[res[c].get(k, 0.0) for k in keys] for c in sorted(res)], dtype=torch.float64, device=dev)
    mn, sm = flat.clone(), flat.clone()
    if world > 1:
        dist.all_reduce(mn, op=dist.ReduceOp.MIN)
        dist.all_reduce(sm, op=dist.ReduceOp.SUM)
    if rank == 0:
        out = {"n_gpus": world, "unit": "GB/s per GPU, all ranks copying at once; min over ranks / mean over ranks", "bytes_per_test": NB}
        for i, c in enumerate(sorted(res)):
            out[c] = {k: [round(float(mn[i, j]), 2), round(float(sm[i, j]) / world, 2)] for j, k in enumerate(keys)}
        out["host"] = {"cpus": os.cpu_count(), "affinity": len(os.sched_getaffinity(0)),
                       "numa": sh("lscpu | grep -i -E 'numa|model name|socket|thread'"),
                       "thp": sh("cat /sys/kernel/mm/transparent_hugepage/enabled"),
                       "pcie": sh("nvidia-smi --query-gpu=index,pcie.link.gen.current,pcie.link.width.current,pcie.link.gen.max --format=csv,noheader"),
                       "topo": sh("nvidia-smi topo -m | head -14"),
                       "iommu": sh("ls /sys/class/iommu 2>/dev/null | head -3; cat /proc/cmdline")}
        print(json.dumps(out, indent=1))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    sys.exit(main())
