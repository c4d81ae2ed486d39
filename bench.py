#!/usr/bin/env python
"""bench.py -- encode+decode throughput of the ntcomp hot path on B200 (BASELINE.json metric).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--config C3]

A step = one pass of the hot path over one batch: encode every read of the rank's shard into ntcomp's
u64 record stream, then decode that stream back to FASTA text.  `value` = whole-job round-trip
Mbases/s (bases / (t_encode + t_decode)) with inputs resident in HBM; `e2e` = the same through the
host-buffer C ABI (ntc_encode_batch + ntc_decode_fasta) with pinned host buffers, H2D/D2H inside
the timed region.  N > 1: one process per GPU (torchrun), reads sharded on 65,536-read boundaries,
index replicated, no data-path collective (weak scaling: every rank gets the config's read count).
At N > 1 the line also carries `c4` (BASELINE config 4: 12 Mbp genome, the 100 M reads split over the N
GPUs -- a strong split) and at N = 8 `c5` (config 5: 2 M x 10 kbp reads at 5 % over 8 GPUs); `value` stays
C3-weak so the driver's scaling efficiency is comparable across N.
`--impl reference` times the CPU oracle's literal restatement of the reference (it cannot be built
here: no Rust toolchain) on a bounded sample with all host threads; that arm loads only oracle/.
"""
import argparse
import gc
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

# BASELINE.json configs (the same table as ntcomp_b200/synth.py; repeated here so that the reference arm does not
# have to import the product package)
CONFIGS = {
    "C2": dict(genome=5_000_000, k=31, n_reads=1_000_000, read_len=150, err_ppm=10_000, seed=20250801 + 2000),
    "C3": dict(genome=5_000_000, k=91, n_reads=10_000_000, read_len=150, err_ppm=10_000, seed=20250801 + 3000),
    "C4": dict(genome=12_000_000, k=91, n_reads=100_000_000, read_len=150, err_ppm=10_000, seed=20250801 + 4000),
    "C5": dict(genome=5_000_000, k=91, n_reads=2_000_000, read_len=10_000, err_ppm=50_000, seed=20250801 + 5000),
}
BLOCK_READS = 65536  # src/main.rs:152

# ms_kernel's HBM streaming bytes per base: 0.25 B of 2-bit packed read in + the
# matching-statistics scratch out (8 B wide entries; 4 B for reads that take the slim path)
MS_HBM_BYTES_PER_BASE = 0.25 + 8.0


def load_peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        try:
            return float(json.load(open(path))["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
        except Exception:
            pass
    return 6650.0, "fallback (B200_PROFILING.md)"


def load_traffic(cfg_name):
    """per-launch ncu counters of the committed `--set full` capture of the newest round (profiles/rN/traffic.json)"""
    try:
        rounds = sorted(d for d in os.listdir(os.path.join(ROOT, "profiles")) if d.startswith("r"))
        for rd in reversed(rounds):
            path = os.path.join(ROOT, "profiles", rd, "traffic.json")
            if os.path.exists(path):
                return json.load(open(path)).get(cfg_name, {}), "profiles/%s/traffic.json" % rd
    except Exception:
        pass
    return {}, None


def config_of(name, reads_override=None):
    cfg = dict(CONFIGS[name])
    if reads_override:
        cfg["n_reads"] = int(reads_override)
    cfg["name"] = name
    return cfg


def shard_bounds(n_reads, world_size, block=BLOCK_READS):
    """the rule of ntcomp_b200/sharding.py (boundaries on 65,536-read multiples)"""
    n_blocks = (n_reads + block - 1) // block
    base, extra = divmod(n_blocks, world_size)
    out, lo_b = [], 0
    for r in range(world_size):
        nb = base + (1 if r < extra else 0)
        out.append((min(n_reads, lo_b * block), min(n_reads, (lo_b + nb) * block)))
        lo_b += nb
    return out


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


def time_oracle(oracle, bases, offsets, threads, budget_s, fast=False, keep_records=False):
    """literal (reference-shaped) encode + decode of the first reads until ~budget_s of CPU work"""
    n_total = len(offsets) - 1
    n = min(n_total, max(threads * 8, 64))
    done_bases, t_enc, t_dec, n_done = 0, 0.0, 0.0, 0
    lo = 0
    kept = []
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
        if keep_records:
            kept.append(rec)
        lo = hi
        if (t_enc + t_dec) < budget_s / 8:
            n *= 2
    return dict(bases=done_bases, reads=n_done, t_enc=t_enc, t_dec=t_dec, records=np.concatenate(kept) if kept else None)


def run_reference(args, rank):
    """--impl reference: the reference's CPU implementation of the path (oracle literal port).  Index, genome and reads
    come from oracle/ alone (its own SBWT builder, the numpy restatement of the generator): the product library is not
    loaded in this arm."""
    if rank != 0:
        return
    from oracle import binding as ob, synth_np
    cfg = config_of(args.config, args.reads)
    threads = os.cpu_count() or 1
    g = synth_np.genome(cfg["genome"])
    oracle = ob.OracleIndex.build([g], cfg["k"])
    n_sample = min(cfg["n_reads"], 1 << 16)
    bases, off = synth_np.reads(g, cfg["seed"], 0, n_sample, cfg["read_len"], cfg["err_ppm"])
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


class Rank:
    """torch / torch.distributed plumbing of this process"""

    def __init__(self):
        import torch
        import torch.distributed as dist
        self.torch, self.dist = torch, dist
        self.rank = int(os.environ.get("RANK", "0"))
        self.local_rank = int(os.environ.get("LOCAL_RANK", "0"))
        self.world = int(os.environ.get("WORLD_SIZE", "1"))
        if not torch.cuda.is_available():
            raise SystemExit("bench.py: no CUDA device; the product path has no CPU fallback (use --impl reference for the CPU arm)")
        torch.cuda.set_device(self.local_rank)
        self.dev = torch.device("cuda", self.local_rank)

    def init(self):
        if self.world > 1:
            self.dist.init_process_group("nccl", device_id=self.dev)

    def barrier(self):
        if self.world > 1:
            self.dist.barrier()
        self.torch.cuda.synchronize()

    def reduce(self, values, op="max"):
        t = self.torch.tensor(values, dtype=self.torch.float64, device=self.dev)
        if self.world > 1:
            self.dist.all_reduce(t, op=self.dist.ReduceOp.MAX if op == "max" else self.dist.ReduceOp.MIN)
        return [float(x) for x in t.tolist()]

    def finish(self):
        if self.world > 1:
            self.dist.destroy_process_group()


def free_host_cache(torch):
    gc.collect()
    torch.cuda.empty_cache()
    try:
        torch._C._host_emptyCache()
    except Exception:
        pass


def measure(R, cfg, lo, n_reads, steps, warm, args, want_e2e=True, e2e_reads_cap=None, want_l2=True, want_columns=False, light=False):
    """One config on this rank's shard [lo, lo + n_reads): device-resident timing (CUDA events on the launching
    stream, barrier + synchronize on both sides, max over ranks) and the end-to-end leg through the host-buffer C ABI.
    light: a secondary config on the N = 1 line -- no host-builder comparison, no pageable-buffer leg."""
    torch = R.torch
    import ctypes as C
    import ntcomp_b200 as nt
    from ntcomp_b200 import _lib, synth
    dev, rank, world = R.dev, R.rank, R.world
    read_len = cfg["read_len"]
    n_bases = n_reads * read_len
    out = {"n_reads": n_reads, "n_bases": n_bases, "lo": lo}

    # ---- index: rank 0 builds and saves, the others load (index replicated per GPU) ----
    t0 = time.time()
    g = synth.genome(cfg["genome"])
    path = "/tmp/ntc_bench_%s_%d_%d.idx" % (cfg["name"], cfg["genome"], cfg["k"])
    index = None
    if rank == 0:
        # `ntcomp build` on the GPU (row f4, ntc_index_build_gpu); at N = 1 with a CPU budget the host builder runs beside it
        # (all host threads) and the two indexes are compared bit for bit
        bstats = {}
        nt.Index.build_gpu([g[:100_000]], cfg["k"], device=R.local_rank)     # CUDA context + module load outside the timing
        tb = time.time()
        index = nt.Index.build_gpu([g], cfg["k"], device=R.local_rank, stats=bstats)
        out["index_build"] = {"impl": "ntc_index_build_gpu", "wall_ms": (time.time() - tb) * 1e3, "device_ms": bstats["device_ms"],
                              "upload_ms": bstats["upload_ms"], "n_sets": index.n_sets, "kmer_occurrences": bstats["occurrences"],
                              "radix_passes": bstats["radix_passes"], "launches": bstats["launches"]}
        if world == 1 and args.cpu_budget > 0 and not light:
            tb = time.time()
            host = nt.Index.build([g], cfg["k"])
            host_s = time.time() - tb
            same = host.n_sets == index.n_sets and (host.lcs() == index.lcs()).all() and all((a == b).all() for a, b in zip(host.rows(), index.rows()))
            if not same:
                raise SystemExit("bench.py: ntc_index_build_gpu and ntc_index_build disagree")
            out["index_build"].update({"host_builder_s": host_s, "host_threads": os.cpu_count(), "identical_to_host_builder": True})
            del host
        if world > 1:
            index.save(path)
    if world > 1:
        R.dist.barrier()
        if rank != 0:
            index = nt.Index.load(path)
    t_index = time.time() - t0
    t0 = time.time()
    ctx = nt.Context(index, R.local_rank)
    # host-buffer leg: both PCIe directions at once by default.  The one-direction-at-a-time mode (NTC_COPY_DIRECTIONS=serial) was
    # built for hosts where mixed traffic is slower than either direction alone (8 GPUs on this pool's boxes) and measured no
    # better there -- ranks drift out of phase, so the host sees mixed traffic anyway (profiles/r2/ab_copy_directions.txt)
    copy_serial = os.environ.get("NTC_COPY_DIRECTIONS", "duplex") == "serial"
    ctx.set_copy_directions(copy_serial)
    out["setup_s"] = {"index_build_or_load": t_index, "ctx_create": time.time() - t0}
    out["index_bytes"] = ctx.index_bytes()

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

    # ---- correctness gate before timing: device round trip must be bit-exact (the record stream itself is compared
    # with the oracle in the cpu_baseline leg below and, at full size, by tests/test_full_size.py) ----
    encode()
    ctx.status(stream)
    n_rec = int(d_cnt[0].item())
    decode(n_rec, fasta=False)
    ctx.status(stream)
    if not (int(d_cnt[0].item()) == n_reads and bool(torch.equal(d_out[:n_bases], d_bases))):
        raise SystemExit("bench.py: decode(encode(reads)) != reads on the device; refusing to time a wrong result")
    out["n_rec"] = n_rec

    # ---- L2 random-sector roofline denominator (measured here) ----
    if want_l2:
        out["l2"] = {"buffer_mb": 32, "gsectors": ctx.l2_random_bench(32 << 20, 256, False),
                     "gsectors_16mb": ctx.l2_random_bench(16 << 20, 256, False), "gsectors_dependent": ctx.l2_random_bench(32 << 20, 256, True)}

    # ---- device-resident timing ----
    for _ in range(warm):
        encode()
        decode(n_rec)
    ctx.status(stream)
    ctx.set_profiling(True)
    ctx.kernel_times(reset=True)
    launches0 = ctx.launch_count()
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(2 * steps + 1)]
    sampler = ClockSampler(R.local_rank)
    R.barrier()
    sampler.start()
    ev[0].record()
    for i in range(steps):
        encode()
        ev[2 * i + 1].record()
        decode(n_rec)
        ev[2 * i + 2].record()
    R.barrier()
    out["clocks"] = sampler.stop()
    ctx.status(stream)
    out["gpu_launches"] = ctx.launch_count() - launches0
    out["ktimes"] = ctx.kernel_times(reset=True)
    ctx.set_profiling(False)
    t_total = ev[0].elapsed_time(ev[2 * steps])
    t_enc = sum(ev[2 * i].elapsed_time(ev[2 * i + 1]) for i in range(steps))
    t_dec = sum(ev[2 * i + 1].elapsed_time(ev[2 * i + 2]) for i in range(steps))
    out["t_total"], out["t_enc"], out["t_dec"] = R.reduce([t_total, t_enc, t_dec])

    # ---- end to end through the host-buffer C ABI, pinned host memory ----
    e2e = None
    if want_e2e:
        ne = n_reads if not e2e_reads_cap else min(n_reads, e2e_reads_cap)
        nbe = ne * read_len
        h_bases = torch.empty(nbe, dtype=torch.uint8).pin_memory()
        h_bases.copy_(d_bases[:nbe])
        h_off = (torch.arange(ne + 1, dtype=torch.int64) * read_len).pin_memory()
        n_rec_e = int(d_roff[ne].item())
        h_rec = torch.empty(n_rec_e + 1024, dtype=torch.int64).pin_memory()
        cap_e = nbe + ne * 24 + 64
        h_text = torch.empty(cap_e, dtype=torch.uint8).pin_memory()
        L = _lib.lib()
        n_rec_c, n_bytes_c, n_seqs_c = C.c_uint64(0), C.c_uint64(0), C.c_uint64(0)
        split = [0.0, 0.0]
        # encode: this rank's share of the host's threads packs the reads to 2 bits per base before they cross PCIe
        # (ntc_ctx_set_host_threads; NTC_HOST_THREADS=0 leaves the packing to the GPU and sends ASCII)
        host_threads = int(os.environ.get("NTC_HOST_THREADS", "-1"))        # library default: host threads only for pageable buffers
        ctx.set_host_threads(host_threads)

        def e2e_step():
            ta = time.perf_counter()
            _lib.check(L.ntc_encode_batch(ctx._h, h_bases.data_ptr(), h_off.data_ptr(), ne, h_rec.data_ptr(), h_rec.numel(), C.byref(n_rec_c), None))
            tb = time.perf_counter()
            _lib.check(L.ntc_decode_fasta(ctx._h, h_rec.data_ptr(), n_rec_c.value, lo + 1, h_text.data_ptr(), cap_e, C.byref(n_bytes_c), C.byref(n_seqs_c)))
            split[0] += tb - ta
            split[1] += time.perf_counter() - tb

        def pcie_peak():
            """pinned-memory copy bandwidth of this box (GB/s), the ceiling of the end-to-end leg"""
            nb = min(nbe, 1 << 30)
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
        R.barrier()                                   # all ranks copy at the same time: the host side is shared
        pcie = R.reduce(pcie_peak(), "min")
        e2e_step()
        first = (b">seq.%d\n" % (lo + 1)) + bytes(h_bases[:read_len].numpy()) + b"\n"
        if n_seqs_c.value != ne or n_rec_c.value != n_rec_e or bytes(h_text[: len(first)].numpy()) != first:
            raise SystemExit("bench.py: end-to-end FASTA does not start with the first read")
        R.barrier()
        split[0] = split[1] = 0.0
        t0 = time.perf_counter()
        for _ in range(e2e_steps):
            e2e_step()
        torch.cuda.synchronize()
        t_e2e = R.reduce([time.perf_counter() - t0])[0]
        h2d = (((nbe + 31) // 32 + 4) * 8 if host_threads > 0 else nbe) + (ne + 1) * 8 + n_rec_e * 8
        d2h = n_rec_e * 8 + int(n_bytes_c.value)
        e2e = {"value": nbe * world * e2e_steps / t_e2e / 1e6, "unit": UNIT, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
               "steps": e2e_steps, "api": "ntc_encode_batch + ntc_decode_fasta, pinned host buffers", "reads_per_gpu": ne,
               "host_pack_threads": host_threads,   # > 0: reads cross PCIe as 0.25 B/base, packed by that many host threads per GPU inside ntc_encode_batch
               "copy_directions": "serial (all input up, then results down)" if copy_serial else "duplex (results travel back while the next chunk goes up)",
               "encode_ms_per_step": 1e3 * split[0] / e2e_steps, "decode_ms_per_step": 1e3 * split[1] / e2e_steps,
               "pcie_h2d_gbs": pcie[0], "pcie_d2h_gbs": pcie[1],   # per GPU, all ranks copying at once (min over ranks)
               "pcie_floor_ms_per_step": (h2d / pcie[0] + d2h / pcie[1]) / 1e6 if pcie[0] else None}
        if world == 1 and not light:
            # the same two calls from PAGEABLE buffers (what a Rust Vec or a numpy array is): the context stages them through its own page-locked
            # buffers with host threads (ntc_ctx_set_host_threads, default auto); reported beside the contract's page-locked figure, not instead of it
            ctx.set_host_threads(-1)
            pg_bases, pg_off = h_bases.numpy().copy(), h_off.numpy().copy()
            pg_rec, pg_text = np.empty(h_rec.numel(), dtype=np.int64), np.empty(cap_e, dtype=np.uint8)
            tp = [1e9, 1e9]
            for _ in range(2):
                ta = time.perf_counter()
                _lib.check(L.ntc_encode_batch(ctx._h, pg_bases.ctypes.data, pg_off.ctypes.data, ne, pg_rec.ctypes.data, len(pg_rec), C.byref(n_rec_c), None))
                tb = time.perf_counter()
                _lib.check(L.ntc_decode_fasta(ctx._h, pg_rec.ctypes.data, n_rec_c.value, lo + 1, pg_text.ctypes.data, cap_e, C.byref(n_bytes_c), C.byref(n_seqs_c)))
                tc = time.perf_counter()
                tp = [min(tp[0], tb - ta), min(tp[1], tc - tb)]
            if bytes(pg_text[: len(first)]) != first or n_seqs_c.value != ne:
                raise SystemExit("bench.py: end-to-end FASTA from pageable buffers does not start with the first read")
            e2e["pageable_buffers"] = {"value": nbe / (tp[0] + tp[1]) / 1e6, "unit": UNIT, "encode_ms": 1e3 * tp[0], "decode_ms": 1e3 * tp[1],
                                       "host_threads": "auto", "note": "best of 2; staged through the context's page-locked buffers by host threads"}
            ctx.set_host_threads(host_threads)
            del pg_bases, pg_off, pg_rec, pg_text
        out["h_sample"] = (h_bases, h_off)
        del h_rec, h_text
        if want_columns:
            out["columns"] = measure_columns(ctx, h_bases, h_off, ne, read_len, lo, load_peaks()[0])
    out["e2e"] = e2e
    out["ctx"], out["index"], out["d_rec"], out["d_roff"] = ctx, index, d_rec, d_roff
    del d_out, d_g
    return out


def measure_columns(ctx, h_bases, h_off, n_reads, read_len, lo, hbm_peak):
    """Row f1: the fused column codec (ntc_encode_columns / ntc_decode_columns_fasta) on the first blocks of the batch.
    Kernel time = CUDA events around the column kernels on the launching stream (classes column_encode / column_decode)."""
    n = min(n_reads, 16 * BLOCK_READS, max(1, (192 << 20) // read_len))          # 16 blocks of short reads, ~200 MB of long ones
    if n < 1:
        return None
    bases = h_bases[: n * read_len].numpy()
    off = h_off[: n + 1].numpy().astype(np.uint64)
    blocks, words, n_rec = ctx.encode_columns(bases, off, BLOCK_READS)          # warm-up (workspaces, pinned buffers)
    kept = [b for b in blocks if b.status == 0]
    ctx.decode_columns_fasta(kept, words, lo + 1, True)
    ctx.set_profiling(True)
    ctx.kernel_times(reset=True)
    reps = 3
    t0 = time.perf_counter()
    for _ in range(reps):
        blocks, words, n_rec = ctx.encode_columns(bases, off, BLOCK_READS)
    t1 = time.perf_counter()
    for _ in range(reps):
        text, n_seqs, n_dec = ctx.decode_columns_fasta(kept, words, lo + 1, True)
    t2 = time.perf_counter()
    kt = ctx.kernel_times(reset=True)
    ctx.set_profiling(False)
    enc_ms, dec_ms = kt["column_encode"][0] / reps, kt["column_decode"][0] / reps
    coded = 8 * len(words)
    first = (b">seq.%d\n" % (lo + 1)) + bytes(bases[:read_len]) + b"\n"
    if n_seqs != sum(int(b.num_reads) for b in kept) or text[: len(first)] != first:
        raise SystemExit("bench.py: fused column decode does not reproduce the reads")
    enc_bytes, dec_bytes = 8 * n_rec + coded, coded + 8 * n_dec
    return {"reads": n, "blocks": len(blocks), "blocks_dropped_q4": len(blocks) - len(kept), "records": n_rec,
            "d2h_bytes_per_record_before": 8.0, "d2h_bytes_per_record_after": coded / n_rec,
            "encode": {"kernel_ms": enc_ms, "launches": kt["column_encode"][1] // reps, "records_per_s": n_rec / (enc_ms * 1e-3),
                       "algorithmic_bytes": enc_bytes, "hbm_achieved_gbs": enc_bytes / (enc_ms * 1e-3) / 1e9,
                       "hbm_frac": enc_bytes / (enc_ms * 1e-3) / 1e9 / hbm_peak,
                       "call_ms_with_encode_and_copies": 1e3 * (t1 - t0) / reps},
            "decode": {"kernel_ms": dec_ms, "launches": kt["column_decode"][1] // reps, "records_per_s": n_dec / (dec_ms * 1e-3),
                       "algorithmic_bytes": dec_bytes, "hbm_achieved_gbs": dec_bytes / (dec_ms * 1e-3) / 1e9,
                       "hbm_frac": dec_bytes / (dec_ms * 1e-3) / 1e9 / hbm_peak,
                       "call_ms_with_walk_and_copies": 1e3 * (t2 - t1) / reps},
            "note": "algorithmic bytes = 8 B per record in + coded words out (encode), the reverse for decode; batches of a few "
                    "blocks are launch / latency bound, not HBM bound"}


def l2_roofline(m, l2_gsect):
    """the roofline that binds ms_kernel: random 32-byte sectors out of L2 / L1TEX (64 B per base)"""
    ms_ms, ms_launches = m["ktimes"]["ms"]
    ms_avg_s = (ms_ms / max(ms_launches, 1)) * 1e-3
    peak = l2_gsect * 32.0
    ach = m["n_bases"] * 64.0 / ms_avg_s / 1e9
    return ms_avg_s, ach, peak


def summarize_extra(m, steps, l2_gsect, note):
    """compact object for a secondary config (c4 / c5) on an N > 1 line"""
    world_bases = m["total_bases"]
    ms_avg_s, ach, peak = l2_roofline(m, l2_gsect)
    return {"workload": note, "reads_per_gpu": m["n_reads"], "value": world_bases * steps / (m["t_total"] * 1e-3) / 1e6, "unit": UNIT,
            "ms_per_step": m["t_total"] / steps,
            "encode_mbases_s": world_bases * steps / (m["t_enc"] * 1e-3) / 1e6, "decode_mbases_s": world_bases * steps / (m["t_dec"] * 1e-3) / 1e6,
            "records_per_read": m["n_rec"] / m["n_reads"], "l2_random": {"achieved": ach, "peak": peak, "unit": "GB/s", "frac": ach / peak},
            "kernel_ms_per_step": {k: v[0] / steps for k, v in m["ktimes"].items()},
            "e2e": m["e2e"], "index_bytes": m["index_bytes"], "clocks": m["clocks"], "gpu_launches": m["gpu_launches"]}


def cpu_baseline_of(m, n_reads, read_len, budget_s):
    """literal oracle on one thread (the reference is single-threaded) over the first reads of the workload, for about budget_s
    seconds; its record stream is compared with the GPU's for the same reads"""
    from oracle import binding as ob
    index = m["index"]
    oracle = ob.OracleIndex.from_arrays(index.k, index.n_sets, index.rows(), index.lcs())
    h_bases, h_off = m["h_sample"]
    n_s = min(n_reads, 1 << 15)
    sb = h_bases[: n_s * read_len].numpy()
    so = h_off[: n_s + 1].numpy().astype(np.uint64)
    r = time_oracle(oracle, sb, so, 1, budget_s, keep_records=True)
    n_chk = int(m["d_roff"][r["reads"]].item())
    gpu_rec = m["d_rec"][:n_chk].cpu().numpy().view(np.uint64)
    if len(r["records"]) != n_chk or not (gpu_rec == r["records"]).all():
        raise SystemExit("bench.py: the GPU record stream differs from the literal oracle's on the first %d reads" % r["reads"])
    return {"value": r["bases"] / (r["t_enc"] + r["t_dec"]) / 1e6, "unit": UNIT, "cores": 1, "kind": "port",
            "sample": "first %d reads of this workload, literal oracle restatement (reference is single-threaded, src/main.rs:162-177)" % r["reads"],
            "encode_mbases_s": r["bases"] / r["t_enc"] / 1e6, "decode_mbases_s": r["bases"] / r["t_dec"] / 1e6,
            "record_stream_equal_to_gpu": True, "records_compared": n_chk}


def release(R, m):
    for key in ("ctx", "index", "d_rec", "d_roff", "h_sample"):
        m.pop(key, None)
    free_host_cache(R.torch)


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
    ap.add_argument("--no-extras", action="store_true", help="skip the objects of the other BASELINE configs (c2 / c4 / c5)")
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference(args, int(os.environ.get("RANK", "0")))
        return

    R = Rank()
    from ntcomp_b200 import sharding
    cpus = sharding.bind_to_gpu_numa(R.local_rank) if R.world > 1 else None
    R.init()
    rank, world = R.rank, R.world
    warm, steps = max(3, args.warmup), max(1, args.steps)
    cfg = config_of(args.config, args.reads)
    if cfg["name"] == "C5" and args.reads is None:
        cfg["n_reads"] = 250_000  # C5 is an 8-GPU config: 2 M reads / 8
    if cfg["name"] == "C4" and args.reads is None:
        cfg["n_reads"] = 191 * 65536  # C4 is 100 M reads over up to 8 GPUs: 12.5 M per GPU, on a 65,536-read boundary
    n_reads, read_len = cfg["n_reads"], cfg["read_len"]

    # ---- this rank's shard: reads [lo, hi) of the global numbering, boundaries on 65,536 ----
    lo, hi = shard_bounds(n_reads * world, world)[rank] if (n_reads % BLOCK_READS == 0) else (rank * n_reads, (rank + 1) * n_reads)
    assert hi - lo == n_reads
    m = measure(R, cfg, lo, n_reads, steps, warm, args, want_e2e=not args.no_e2e, want_columns=True)
    n_bases, n_rec, ktimes, t_total = m["n_bases"], m["n_rec"], m["ktimes"], m["t_total"]
    l2 = m["l2"]

    # ---- CPU baseline beside it (rank 0, N = 1 only): literal oracle, one thread like the reference; its records
    # double as a parity check of the GPU's record stream for the same reads ----
    cpu = None
    if world == 1 and rank == 0 and args.cpu_budget > 0 and m.get("h_sample") is not None:
        cpu = cpu_baseline_of(m, n_reads, read_len, args.cpu_budget)
    release(R, m)

    # ---- N > 1: BASELINE configs 4 and 5 beside the C3-weak headline ----
    extras = {}
    if world > 1 and cfg["name"] == "C3" and args.reads is None and not args.no_extras:
        c4 = config_of("C4")
        lo4, hi4 = shard_bounds(c4["n_reads"], world)[rank]
        # a GPU's share is worked through in passes of at most 382 blocks (25.0 M reads, ~70 GB of workspaces): 2 passes at N = 2
        passes = shard_bounds(hi4 - lo4, -(-(hi4 - lo4) // (382 * BLOCK_READS)))
        m4 = None
        for (a, b) in passes:
            mp = measure(R, c4, lo4 + a, b - a, 3, 3, args, want_e2e=(m4 is None and not args.no_e2e), e2e_reads_cap=191 * 65536, want_l2=False)
            release(R, mp)
            if m4 is None:
                m4 = mp
            else:
                for key in ("t_total", "t_enc", "t_dec", "n_rec", "n_reads", "gpu_launches"):
                    m4[key] += mp[key]
                m4["ktimes"] = {k: (m4["ktimes"][k][0] + mp["ktimes"][k][0], m4["ktimes"][k][1] + mp["ktimes"][k][1]) for k in m4["ktimes"]}
        m4["total_bases"] = c4["n_reads"] * c4["read_len"]
        extras["c4"] = summarize_extra(m4, 3, l2["gsectors"], "C4: 12 Mbp genome, k=91, 100 M x 150 bp reads split over %d GPUs on 65,536-read "
                                       "boundaries (strong split, %d pass(es) of <= 25.0 M reads per GPU; e2e leg on the first <= 12.5 M "
                                       "reads of each shard)" % (world, len(passes)))
        release(R, m4)
        if world == 8:
            c5 = config_of("C5")
            n5 = c5["n_reads"] // world
            m5 = measure(R, c5, rank * n5, n5, 3, 3, args, want_e2e=not args.no_e2e, want_l2=False)
            m5["total_bases"] = n5 * world * c5["read_len"]
            extras["c5"] = summarize_extra(m5, 3, l2["gsectors"], "C5: 5 Mbp genome, k=91, 2 M x 10 kbp reads at 5 % over 8 GPUs")
            release(R, m5)

    # ---- N = 1: the other BASELINE configs beside the C3 headline -- C2 whole (it is a one-B200 config), and one 8-GPU shard each
    # of C4 and C5 (what one GPU of the N = 8 line works on), so that every config has a driver-run number on this line ----
    if world == 1 and cfg["name"] == "C3" and args.reads is None and not args.no_extras:
        c2 = config_of("C2")
        m2 = measure(R, c2, 0, c2["n_reads"], 3, 3, args, want_e2e=not args.no_e2e, want_l2=False, light=True)
        m2["total_bases"] = c2["n_reads"] * c2["read_len"]
        extras["c2"] = summarize_extra(m2, 3, l2["gsectors"], "C2: 5 Mbp genome, k=31, 1 M x 150 bp reads at 1 %, whole config on 1 GPU")
        if args.cpu_budget > 0 and m2.get("h_sample") is not None:
            extras["c2"]["cpu_baseline"] = cpu_baseline_of(m2, c2["n_reads"], c2["read_len"], min(args.cpu_budget, 5.0))
        release(R, m2)
        c4 = config_of("C4")
        lo4, hi4 = shard_bounds(c4["n_reads"], 8)[0]
        m4 = measure(R, c4, lo4, hi4 - lo4, 3, 3, args, want_e2e=not args.no_e2e, want_l2=False, light=True)
        m4["total_bases"] = (hi4 - lo4) * c4["read_len"]
        extras["c4"] = summarize_extra(m4, 3, l2["gsectors"], "C4 shard: 12 Mbp genome, k=91 (64-node rank blocks), shard 0 of the 8-GPU split of 100 M x "
                                       "150 bp reads (%d reads) on 1 GPU" % (hi4 - lo4))
        release(R, m4)
        c5 = config_of("C5")
        n5 = c5["n_reads"] // 8
        m5 = measure(R, c5, 0, n5, 3, 3, args, want_e2e=not args.no_e2e, want_l2=False, light=True)
        m5["total_bases"] = n5 * c5["read_len"]
        extras["c5"] = summarize_extra(m5, 3, l2["gsectors"], "C5 shard: 5 Mbp genome, k=91, shard 0 of the 8-GPU split of 2 M x 10 kbp reads at 5 %% "
                                       "(%d reads) on 1 GPU" % n5)
        release(R, m5)

    if rank != 0:
        R.finish()
        return

    total_bases = n_bases * world
    value = total_bases * steps / (t_total * 1e-3) / 1e6
    hbm_peak, peak_src = load_peaks()
    traffic, traffic_src = load_traffic(cfg["name"]) if args.reads is None else ({}, None)
    ms_cnt = traffic.get("ms_kernel") if isinstance(traffic.get("ms_kernel"), dict) else {"dram_bytes": traffic.get("ms_kernel")}
    walk_cnt = traffic.get("walk_kernel") if isinstance(traffic.get("walk_kernel"), dict) else {"dram_bytes": traffic.get("walk_kernel")}
    ms_avg_s, l2_ach, l2_peak = l2_roofline(m, l2["gsectors"])
    walk_ms, walk_launches = ktimes["walk"]
    walk_avg_s = (walk_ms / max(walk_launches, 1)) * 1e-3
    top = max(ktimes.items(), key=lambda kv: kv[1][0])
    rec_per_base = n_rec / n_bases
    sect = (ms_cnt.get("load_sectors", 0) or 0) + (ms_cnt.get("store_sectors", 0) or 0)
    # SURVEY 8(d) / DESIGN.md section 4: one ms_kernel launch processes n_bases bases.  The roofline that binds is random
    # 32-byte sectors out of L2 / L1TEX -- 64 B per base (2 rank probes x 32 B) against the rate ntc_l2_random_bench
    # measures on this GPU in this run.  The kernel's HBM streaming (0.25 B/base of packed reads in + the scratch out)
    # is reported beside it as `hbm`, with the ncu-counted DRAM traffic.
    roofline = {
        "kernel": "ms_kernel (k-bounded matching statistics)", "bound": "l2_random_sectors",
        "achieved": l2_ach, "peak": l2_peak, "unit": "GB/s", "frac": l2_ach / l2_peak,
        "algorithmic_bytes_per_base": 64, "algorithmic_bytes_per_launch": n_bases * 64,
        "peak_source": "ntc_l2_random_bench in this run: independent 16 B loads at hashed addresses of a %d MB buffer, %.1f G sectors/s "
                       "(16 MB: %.1f, dependent chain: %.1f); 148 SMs x 1 sector / clk" % (l2["buffer_mb"], l2["gsectors"], l2["gsectors_16mb"],
                                                                                        l2["gsectors_dependent"]),
        "traffic": ms_cnt.get("dram_bytes"), "traffic_source": traffic_src,
        "sectors_moved": {"per_base": sect / n_bases if sect else None, "frac_of_peak": sect / ms_avg_s / 1e9 / l2["gsectors"] if sect else None,
                          "source": traffic_src, "note": "ncu-counted L1TEX load + store sectors of one launch over this run's launch time"},
        "hbm": {"bound": "hbm", "algorithmic_bytes_per_base": MS_HBM_BYTES_PER_BASE, "achieved": n_bases * MS_HBM_BYTES_PER_BASE / ms_avg_s / 1e9,
                "peak": hbm_peak, "unit": "GB/s", "frac": n_bases * MS_HBM_BYTES_PER_BASE / ms_avg_s / 1e9 / hbm_peak, "peak_source": peak_src,
                "note": "0.25 B/base of 2-bit packed reads in + matching-statistics scratch out; the kernel is not HBM bound"},
        "kernel_ms_per_launch": {k: (v[0] / v[1] if v[1] else 0.0) for k, v in ktimes.items()},
        "kernel_ms_per_step": {k: v[0] / steps for k, v in ktimes.items()},
        "kernel_share_of_step": {k: v[0] / (t_total if t_total else 1) for k, v in ktimes.items()},
        "dominant": top[0],
        # walk_kernel: one 8-byte jump-table entry (one sector) per 16 bases = 2 B/base of random sectors; HBM streaming =
        # 8 B x records/base in + 1 B/base + FASTA framing out
        "decode_walk": {"l2_random": {"algorithmic_bytes_per_base": 2.0, "achieved": n_bases * 2.0 / walk_avg_s / 1e9, "peak": l2_peak, "unit": "GB/s",
                                      "frac": n_bases * 2.0 / walk_avg_s / 1e9 / l2_peak},
                        "hbm": {"algorithmic_bytes_per_base": 8.0 * rec_per_base + 1.08, "achieved": n_bases * (8.0 * rec_per_base + 1.08) / walk_avg_s / 1e9,
                                "peak": hbm_peak, "unit": "GB/s", "frac": n_bases * (8.0 * rec_per_base + 1.08) / walk_avg_s / 1e9 / hbm_peak,
                                "traffic": walk_cnt.get("dram_bytes")}},
    }
    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": steps, "warmup": warm, "ms_per_step": t_total / steps,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u64/u8 integer", "data": "synthetic",
        "config": workload_config(cfg, world),
        "encode_mbases_s": total_bases * steps / (m["t_enc"] * 1e-3) / 1e6, "decode_mbases_s": total_bases * steps / (m["t_dec"] * 1e-3) / 1e6,
        "records_per_read": n_rec / n_reads, "clocks": m["clocks"],
        "e2e": m["e2e"],
        "host_binding": ("NUMA-local CPUs of each GPU (%d CPUs on rank 0)" % len(cpus)) if cpus else "none",
        "gpu_launches": m["gpu_launches"], "roofline": roofline, "cpu_baseline": cpu, "setup_s": m["setup_s"],
        "index_bytes": m["index_bytes"], "index_build": m.get("index_build"), "columns": m.get("columns"),
    }
    line.update(extras)
    emit(line)
    R.finish()


if __name__ == "__main__":
    main()
