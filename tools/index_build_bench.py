"""Row f4 measurement: ntc_index_build_gpu against the host builder (ntc_index_build, all host threads) on the BASELINE
genomes; checks the two indexes are identical and writes one JSON object per line."""
import json
import sys
import time

import numpy as np

sys.path.insert(0, __file__.rsplit("/", 2)[0])
import ntcomp_b200 as nt  # noqa: E402
from ntcomp_b200 import synth  # noqa: E402


def main():
    out = open(sys.argv[1], "w") if len(sys.argv) > 1 else sys.stdout
    nt.Index.build_gpu([synth.genome(50_000)], 31)            # CUDA context + module load outside the timings
    for G, k in ((5_000_000, 31), (5_000_000, 91), (12_000_000, 91), (5_000_000, 255)):
        g = synth.genome(G)
        t = time.time()
        host = nt.Index.build([g], k)
        host_s = time.time() - t
        runs = []
        for _ in range(3):
            st = {}
            t = time.time()
            gpu = nt.Index.build_gpu([g], k, stats=st)
            st["wall_ms"] = (time.time() - t) * 1e3
            runs.append(st)
        same = gpu.n_sets == host.n_sets and (gpu.lcs() == host.lcs()).all() and all((a == b).all() for a, b in zip(gpu.rows(), host.rows()))
        best = min(runs, key=lambda r: r["wall_ms"])
        line = {"genome_bp": G, "k": k, "n_sets": int(host.n_sets), "identical": bool(same), "host_builder_s": round(host_s, 3),
                "host_threads": int(__import__("os").cpu_count()), "gpu_wall_ms": round(best["wall_ms"], 2),
                "gpu_device_ms": round(best["device_ms"], 2), "gpu_upload_ms": round(best["upload_ms"], 2),
                "radix_passes": best["radix_passes"], "launches": best["launches"], "occurrences": best["occurrences"],
                "speedup_vs_host": round(host_s * 1e3 / best["wall_ms"], 1)}
        out.write(json.dumps(line) + "\n")
        out.flush()


if __name__ == "__main__":
    main()
