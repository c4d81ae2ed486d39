#!/usr/bin/env python
"""bench.py -- encode+decode throughput of the ntcomp hot path on B200 (BASELINE.json metric).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--config C3]

A step = one pass of the hot path over one batch: encode every read of the rank's shard into ntcomp's
u64 record stream, then decode that stream back to FASTA text.  `value` = whole-job round-trip
Mbases/s (bases / (t_encode + t_decode)) with inputs resident in HBM; `e2e` = the same through the
host-buffer C ABI (ntc_encode_batch + ntc_decode_fasta) with pinned host buffers, H2D/D2H inside
the timed region.  N > 1: one process per GPU (torchrun), reads sharded on 65,536-read boundaries,
index replicated, no data-path collective (weak scaling: every rank gets the config's read count).
`--impl reference` times the CPU oracle's literal restatement of the reference (it cannot be built
here: no Rust toolchain) on a bounded sample with all host threads.
"""
import argparse
import json
import os
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

import numpy as np  # noqa: E402

METRIC = "encode+decode round-trip throughput"
UNIT = "Mbases/s"


def load_peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        try:
            return float(json.load(open(path))["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
        except Exception:
            pass
    return 6650.0, "fallback (B200_PROFILING.md)"


def load_traffic(cfg_name, kernel):
    """DRAM bytes per launch of `kernel` from the committed ncu capture (profiles/), or None"""
    try:
        rounds = sorted(d for d in os.listdir(os.path.join(ROOT, "profiles")) if d.startswith("r"))
        for rd in reversed(rounds):
            path = os.path.join(ROOT, "profiles", rd, "traffic.json")
            if os.path.exists(path):
                return json.load(open(path)).get(cfg_name, {}).get(kernel)
    except Exception:
        pass
    return None


def config_of(name, reads_override=None):
    from ntcomp_b200 import synth
    cfg = dict(synth.CONFIGS[name])
    if reads_override:
        cfg["n_reads"] = int(reads_override)
    cfg["name"] = name
    return cfg


class ClockSampler(threading.Thread):
    """samples SM clock / throttle reasons of one GPU while the timed region runs"""

    def __init__(self, device):
        super().__init__(daemon=True)
        self.device, self.samples, self.reasons, self.max_mhz, self._halt = device, [], set(), None, threading.Event()
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(device)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
        except Exception:
            self.nv = None

    def run(self):
        if self.nv is None:
            return
        nv = self.nv
        names = {nv.nvmlClocksThrottleReasonHwSlowdown: "hw_slowdown", nv.nvmlClocksThrottleReasonHwThermalSlowdown: "hw_thermal_slowdown",
                 nv.nvmlClocksThrottleReasonSwThermalSlowdown: "sw_thermal_slowdown", nv.nvmlClocksThrottleReasonSwPowerCap: "sw_power_cap"}
        while not self._halt.is_set():
            try:
                self.samples.append(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM))
                mask = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                for bit, nm in names.items():
                    if mask & bit:
                        self.reasons.add(nm)
            except Exception:
                pass
            time.sleep(0.005)

    def stop(self):
        self._halt.set()
        self.join(timeout=2)
        return {"sm_mhz": float(np.median(self.samples)) if self.samples else None, "sm_max_mhz": self.max_mhz,
                "reasons": sorted(self.reasons), "samples": len(self.samples)}


def oracle_from(index):
    from oracle import binding as ob
    return ob.OracleIndex.from_arrays(index.k, index.n_sets, index.rows(), index.lcs())


def time_oracle(oracle, bases, offsets, threads, budget_s, fast=False):
    """literal (reference-shaped) encode + decode of the first reads until ~budget_s of CPU work"""
    n_total = len(offsets) - 1
    n = min(n_total, max(threads * 8, 64))
    done_bases, t_enc, t_dec, n_done = 0, 0.0, 0.0, 0
    lo = 0
    while lo < n_total and (t_enc + t_dec) < budget_s:
        hi = min(n_total, lo + n)
        sub_off = offsets[lo:hi + 1]
        sub = bases[int(sub_off[0]): int(sub_off[-1])]
        t0 = time.perf_counter()
        rec, _ = oracle.encode_reads(sub, sub_off - sub_off[0], fast=fast, threads=threads)
        t1 = time.perf_counter()
        if threads == 1:
            oracle.decode(rec, fast=fast)
        else:  # decode_sequence is per block; spread blocks over threads
            from concurrent.futures import ThreadPoolExecutor
            roff = np.searchsorted(np.cumsum(((rec >> np.uint64(56)) & np.uint64(1)).astype(np.int64)), np.linspace(0, hi - lo, threads + 1)[1:-1].astype(np.int64) + 1)
            parts = np.split(rec, roff)
            with ThreadPoolExecutor(threads) as ex:
                list(ex.map(lambda p: oracle.decode(p, fast=fast) if len(p) else None, parts))
        t2 = time.perf_counter()
        t_enc += t1 - t0
        t_dec += t2 - t1
        done_bases += int(sub_off[-1] - sub_off[0])
        n_done += hi - lo
        lo = hi
        if (t_enc + t_dec) < budget_s / 8:
            n *= 2
    return dict(bases=done_bases, reads=n_done, t_enc=t_enc, t_dec=t_dec)


def run_reference(args, rank):
    """--impl reference: the reference's CPU implementation of the path (oracle literal port)."""
    if rank != 0:
        return
    import ntcomp_b200 as nt
    from ntcomp_b200 import synth
    cfg = config_of(args.config, args.reads)
    threads = os.cpu_count() or 1
    g = synth.genome(cfg["genome"])
    index = nt.Index.build([g], cfg["k"])
    oracle = oracle_from(index)
    n_sample = min(cfg["n_reads"], 1 << 16)
    bases, off = synth.reads(g, cfg["seed"], 0, n_sample, cfg["read_len"], cfg["err_ppm"])
    steps, warm = max(1, args.steps), max(0, args.warmup)
    budget = max(2.0, min(20.0, 120.0 / (steps + warm)))
    results = []
    for i in range(warm + steps):
        r = time_oracle(oracle, bases, off, threads, budget)
        if i >= warm:
            results.append(r)
    tb = sum(r["bases"] for r in results)
    tt = sum(r["t_enc"] + r["t_dec"] for r in results)
    value = tb / tt / 1e6
    sample = "first %d reads of %s per step (literal oracle: 4 x k-char search left extension, k selects per access_kmer), %d threads" % (
        results[-1]["reads"], cfg["name"], threads)
    line = {"impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": steps, "warmup": warm,
            "ms_per_step": 1e3 * tt / steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u64/u8 integer",
            "data": "synthetic", "config": workload_config(cfg, args.gpus),
            "encode_mbases_s": tb / sum(r["t_enc"] for r in results) / 1e6, "decode_mbases_s": tb / sum(r["t_dec"] for r in results) / 1e6,
            "cpu_baseline": {"value": value, "unit": UNIT, "cores": threads, "kind": "port", "sample": sample},
            "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}, "gpu_launches": 0}
    emit(line)


def workload_config(cfg, n_gpus):
    return {"workload": "%s: synthetic %d bp genome (splitmix64, both strands indexed), k=%d, %d x %d bp reads per GPU at %.1f%% substitutions, "
                        "encode to u64 records then decode to FASTA" % (cfg["name"], cfg["genome"], cfg["k"], cfg["n_reads"], cfg["read_len"],
                                                                        cfg["err_ppm"] / 1e4),
            "reads_per_gpu": cfg["n_reads"], "read_len": cfg["read_len"], "k": cfg["k"], "genome_bp": cfg["genome"],
            "parallelism": "reads sharded x%d, index replicated, no collective" % n_gpus,
            "l2_policy": "inputs larger than L2 (%.0f MB of reads per step per GPU)" % (cfg["n_reads"] * cfg["read_len"] / 1e6)}


_REAL_STDOUT = None


def emit(line):
    """the ONE JSON line, on the process's original stdout"""
    data = (json.dumps(line) + "\n").encode()
    if _REAL_STDOUT is None:
        sys.stdout.write(data.decode())
        sys.stdout.flush()
    else:
        os.write(_REAL_STDOUT, data)


def main():
    # libraries print to fd 1 behind Python's back (NCCL's version banner when NCCL_DEBUG is set): route fd 1 to
    # stderr for the run and keep the original for the result line
    global _REAL_STDOUT
    sys.stdout.flush()
    _REAL_STDOUT = os.dup(1)
    os.dup2(2, 1)
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--config", default="C3", choices=["C2", "C3", "C4", "C5"])
    ap.add_argument("--reads", type=int, default=None, help="override reads per GPU (testing)")
    ap.add_argument("--cpu-budget", type=float, default=15.0, help="seconds of CPU baseline work (0 = skip)")
    ap.add_argument("--no-e2e", action="store_true", help="skip the host-buffer end-to-end leg (profiling runs)")
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if args.impl == "reference":
        run_reference(args, rank)
        return

    import torch
    import torch.distributed as dist
    import ntcomp_b200 as nt
    from ntcomp_b200 import _lib, sharding, synth

    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the product path has no CPU fallback (use --impl reference for the CPU arm)")
    cpus = sharding.bind_to_gpu_numa(local_rank) if world > 1 else None
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    warm, steps = max(3, args.warmup), max(1, args.steps)
    cfg = config_of(args.config, args.reads)
    if cfg["name"] == "C5" and args.reads is None:
        cfg["n_reads"] = 250_000  # C5 is an 8-GPU config: 2 M reads / 8
    if cfg["name"] == "C4" and args.reads is None:
        cfg["n_reads"] = 191 * 65536  # C4 is 100 M reads over up to 8 GPUs: 12.5 M per GPU, on a 65,536-read boundary
    n_reads, read_len = cfg["n_reads"], cfg["read_len"]
    n_bases = n_reads * read_len

    # ---- index: rank 0 builds and saves, the others load (index replicated per GPU) ----
    t0 = time.time()
    g = synth.genome(cfg["genome"])
    path = "/tmp/ntc_bench_%s_%d_%d.idx" % (cfg["name"], cfg["genome"], cfg["k"])
    if rank == 0:
        index = nt.Index.build([g], cfg["k"])
        if world > 1:
            index.save(path)
    if world > 1:
        dist.barrier()
        if rank != 0:
            index = nt.Index.load(path)
    ctx = nt.Context(index, local_rank)
    t_setup = time.time() - t0

    # ---- this rank's shard: reads [lo, hi) of the global numbering, boundaries on 65,536 ----
    lo, hi = sharding.shard_bounds(n_reads * world, world)[rank] if (n_reads % sharding.BLOCK_READS == 0) else (rank * n_reads, (rank + 1) * n_reads)
    assert hi - lo == n_reads
    stream = torch.cuda.current_stream().cuda_stream
    d_g = torch.from_numpy(g).to(dev)
    d_bases = torch.empty(n_bases, dtype=torch.uint8, device=dev)
    _lib.check(_lib.lib().ntc_synth_reads_device(ctx._h, d_g.data_ptr(), len(g), cfg["seed"], lo, n_reads, read_len, cfg["err_ppm"], d_bases.data_ptr(), stream))
    d_off = torch.arange(n_reads + 1, dtype=torch.int64, device=dev) * read_len
    cap = max(n_bases // 6, 1 << 20)
    d_rec = torch.empty(cap, dtype=torch.int64, device=dev)
    d_roff = torch.empty(n_reads + 1, dtype=torch.int64, device=dev)
    d_cnt = torch.zeros(4, dtype=torch.int64, device=dev)
    out_cap = n_bases + n_reads * 24 + 64
    d_out = torch.empty(out_cap, dtype=torch.uint8, device=dev)

    def encode():
        ctx.encode_device(d_bases.data_ptr(), d_off.data_ptr(), n_reads, n_bases, d_rec.data_ptr(), cap, d_roff.data_ptr(), d_cnt.data_ptr(), stream)

    def decode(n_rec, fasta=True):
        ctx.decode_device(d_rec.data_ptr(), n_rec, lo + 1, fasta, d_out.data_ptr(), out_cap, 0, 0, d_cnt.data_ptr(), stream)

    # ---- correctness gate before timing: device round trip must be bit-exact ----
    encode()
    ctx.status(stream)
    n_rec = int(d_cnt[0].item())
    decode(n_rec, fasta=False)
    ctx.status(stream)
    if not (int(d_cnt[0].item()) == n_reads and bool(torch.equal(d_out[:n_bases], d_bases))):
        raise SystemExit("bench.py: decode(encode(reads)) != reads on the device; refusing to time a wrong result")

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # ---- L2 random-sector roofline denominator (measured here) ----
    l2_gsect = ctx.l2_random_bench(48 << 20, 256, False)
    l2_gsect_small = ctx.l2_random_bench(16 << 20, 256, False)
    l2_dep = ctx.l2_random_bench(48 << 20, 256, True)

    # ---- device-resident timing ----
    for _ in range(warm):
        encode()
        decode(n_rec)
    ctx.status(stream)
    ctx.set_profiling(True)
    ctx.kernel_times(reset=True)
    launches0 = ctx.launch_count()
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(2 * steps + 1)]
    sampler = ClockSampler(local_rank)
    barrier()
    sampler.start()
    ev[0].record()
    for i in range(steps):
        encode()
        ev[2 * i + 1].record()
        decode(n_rec)
        ev[2 * i + 2].record()
    barrier()
    clocks = sampler.stop()
    ctx.status(stream)
    gpu_launches = ctx.launch_count() - launches0
    ktimes = ctx.kernel_times(reset=True)
    ctx.set_profiling(False)
    t_total = ev[0].elapsed_time(ev[2 * steps])
    t_enc = sum(ev[2 * i].elapsed_time(ev[2 * i + 1]) for i in range(steps))
    t_dec = sum(ev[2 * i + 1].elapsed_time(ev[2 * i + 2]) for i in range(steps))
    tt = torch.tensor([t_total, t_enc, t_dec], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
    t_total, t_enc, t_dec = [float(x) for x in tt.tolist()]

    # ---- end to end through the host-buffer C ABI, pinned host memory ----
    h_bases = torch.empty(n_bases, dtype=torch.uint8).pin_memory()
    h_bases.copy_(d_bases)
    h_off = (torch.arange(n_reads + 1, dtype=torch.int64) * read_len).pin_memory()
    h_rec = torch.empty(n_rec + 1024, dtype=torch.int64).pin_memory()
    h_text = torch.empty(out_cap, dtype=torch.uint8).pin_memory()
    import ctypes as C
    L = _lib.lib()
    n_rec_c, n_bytes_c, n_seqs_c = C.c_uint64(0), C.c_uint64(0), C.c_uint64(0)

    e2e_split = [0.0, 0.0]

    def e2e_step():
        ta = time.perf_counter()
        _lib.check(L.ntc_encode_batch(ctx._h, h_bases.data_ptr(), h_off.data_ptr(), n_reads, h_rec.data_ptr(), h_rec.numel(), C.byref(n_rec_c), None))
        tb = time.perf_counter()
        _lib.check(L.ntc_decode_fasta(ctx._h, h_rec.data_ptr(), n_rec_c.value, lo + 1, h_text.data_ptr(), out_cap, C.byref(n_bytes_c), C.byref(n_seqs_c)))
        e2e_split[0] += tb - ta
        e2e_split[1] += time.perf_counter() - tb

    def pcie_peak():
        """pinned-memory copy bandwidth of this box (GB/s), the ceiling of the end-to-end leg"""
        nb = min(n_bases, 1 << 30)
        e0, e1, e2 = (torch.cuda.Event(enable_timing=True) for _ in range(3))
        best = [0.0, 0.0]
        for _ in range(3):
            e0.record()
            d_out[:nb].copy_(h_text[:nb], non_blocking=True)
            e1.record()
            h_text[:nb].copy_(d_out[:nb], non_blocking=True)
            e2.record()
            torch.cuda.synchronize()
            best = [max(best[0], nb / e0.elapsed_time(e1) / 1e6), max(best[1], nb / e1.elapsed_time(e2) / 1e6)]
        return best

    e2e_steps = max(1, min(steps, 3))
    t_e2e = float("nan")
    pcie = [None, None]
    if not args.no_e2e:
        barrier()                                   # all ranks copy at the same time: the host side is shared
        pcie = pcie_peak()
        if world > 1:
            tp = torch.tensor(pcie, dtype=torch.float64, device=dev)
            dist.all_reduce(tp, op=dist.ReduceOp.MIN)
            pcie = [float(x) for x in tp.tolist()]
        e2e_step()
        first = (b">seq.%d\n" % (lo + 1)) + bytes(h_bases[:read_len].numpy()) + b"\n"
        if n_seqs_c.value != n_reads or bytes(h_text[: len(first)].numpy()) != first:
            raise SystemExit("bench.py: end-to-end FASTA does not start with the first read")
        barrier()
        e2e_split[0] = e2e_split[1] = 0.0
        t0 = time.perf_counter()
        for _ in range(e2e_steps):
            e2e_step()
        torch.cuda.synchronize()
        t_e2e = time.perf_counter() - t0
    te = torch.tensor([t_e2e], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(te, op=dist.ReduceOp.MAX)
    t_e2e = float(te.item())
    h2d = n_bases + (n_reads + 1) * 8 + n_rec * 8
    d2h = n_rec * 8 + int(n_bytes_c.value)

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return

    total_bases = n_bases * world
    value = total_bases * steps / (t_total * 1e-3) / 1e6
    hbm_peak, peak_src = load_peaks()
    ms_ms, ms_launches = ktimes["ms"]
    walk_ms, walk_launches = ktimes["walk"]
    top = max(ktimes.items(), key=lambda kv: kv[1][0])
    # SURVEY §8(d) / DESIGN.md §4: one ms_kernel launch processes n_bases bases.  Random L2 sectors: 64 B/base
    # (2 rank probes x 32 B).  HBM streaming of this kernel: 0.25 B/base of 2-bit packed reads in + 8 B/base of
    # matching-statistics scratch out (the ASCII byte in and the records out belong to pack2bit / emit).
    ms_avg_s = (ms_ms / max(ms_launches, 1)) * 1e-3
    l2_peak_gbs = l2_gsect * 32.0
    hbm_bytes_per_base = 0.25 + 8.0
    roofline = {
        "kernel": "ms_kernel (k-bounded matching statistics)", "bound": "hbm",
        "achieved": n_bases * hbm_bytes_per_base / ms_avg_s / 1e9, "peak": hbm_peak, "unit": "GB/s",
        "frac": n_bases * hbm_bytes_per_base / ms_avg_s / 1e9 / hbm_peak, "peak_source": peak_src,
        "algorithmic_bytes_per_base": hbm_bytes_per_base, "algorithmic_bytes_per_launch": n_bases * hbm_bytes_per_base,
        "traffic": load_traffic(cfg["name"], "ms_kernel") if args.reads is None else None,
        "binding": "l2_random",
        "note": "the kernel is bound by random 32 B sectors out of L2 / L1TEX (index is L2 resident), not by HBM: l2_random is the roofline that "
                "binds (SURVEY 8d, primary); traffic above the algorithmic bytes = the record-count pass re-reading scratch + index lines "
                "evicted by the scratch stream (DESIGN.md section 4)",
        "l2_random": {"algorithmic_bytes_per_base": 64, "achieved": n_bases * 64.0 / ms_avg_s / 1e9, "peak": l2_peak_gbs, "unit": "GB/s",
                      "frac": n_bases * 64.0 / ms_avg_s / 1e9 / l2_peak_gbs,
                      "peak_source": "ntc_l2_random_bench: independent 16 B loads, 48 MB buffer, %.1f G sectors/s (16 MB: %.1f, dependent chain: %.1f)" % (
                          l2_gsect, l2_gsect_small, l2_dep)},
        "kernel_ms_per_launch": {k: (v[0] / v[1] if v[1] else 0.0) for k, v in ktimes.items()},
        "kernel_ms_per_step": {k: v[0] / steps for k, v in ktimes.items()},
        "kernel_share_of_step": {k: v[0] / (t_total if t_total else 1) for k, v in ktimes.items()},
        "dominant": top[0],
        "decode_walk": {"algorithmic_bytes_per_base": 32, "achieved": n_bases * 32.0 / ((walk_ms / max(walk_launches, 1)) * 1e-3) / 1e9,
                        "peak": l2_peak_gbs, "unit": "GB/s"},
    }
    # ---- CPU baseline beside it (rank 0, N = 1 only): literal oracle, one thread like the reference ----
    cpu = None
    if world == 1 and args.cpu_budget > 0:
        oracle = oracle_from(index)
        n_s = min(n_reads, 1 << 15)
        sb = h_bases[: n_s * read_len].numpy()
        so = h_off[: n_s + 1].numpy().astype(np.uint64)
        r = time_oracle(oracle, sb, so, 1, args.cpu_budget)
        cpu = {"value": r["bases"] / (r["t_enc"] + r["t_dec"]) / 1e6, "unit": UNIT, "cores": 1, "kind": "port",
               "sample": "first %d reads of this workload, literal oracle restatement (reference is single-threaded, src/main.rs:162-177)" % r["reads"],
               "encode_mbases_s": r["bases"] / r["t_enc"] / 1e6, "decode_mbases_s": r["bases"] / r["t_dec"] / 1e6}
    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": steps, "warmup": warm, "ms_per_step": t_total / steps,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u64/u8 integer", "data": "synthetic",
        "config": workload_config(cfg, world),
        "encode_mbases_s": total_bases * steps / (t_enc * 1e-3) / 1e6, "decode_mbases_s": total_bases * steps / (t_dec * 1e-3) / 1e6,
        "records_per_read": n_rec / n_reads, "clocks": clocks,
        "e2e": {"value": total_bases * e2e_steps / t_e2e / 1e6, "unit": UNIT, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                "steps": e2e_steps, "api": "ntc_encode_batch + ntc_decode_fasta, pinned host buffers",
                "encode_ms_per_step": 1e3 * e2e_split[0] / e2e_steps, "decode_ms_per_step": 1e3 * e2e_split[1] / e2e_steps,
                "pcie_h2d_gbs": pcie[0], "pcie_d2h_gbs": pcie[1],   # per GPU, all ranks copying at once (min over ranks)
                "pcie_floor_ms_per_step": (h2d / pcie[0] + d2h / pcie[1]) / 1e6 if pcie[0] else None},
        "host_binding": ("NUMA-local CPUs of each GPU (%d CPUs on rank 0)" % len(cpus)) if cpus else "none",
        "gpu_launches": gpu_launches, "roofline": roofline, "cpu_baseline": cpu, "setup_s": t_setup,
        "index_bytes": ctx.index_bytes(),
    }
    emit(line)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
