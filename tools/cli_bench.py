#!/usr/bin/env python
"""File-to-file throughput of the command-line front end (row f3): synthetic reads of a BASELINE config
written as FASTA, `ntcomp-b200 build / encode / decode`, wall-clock Mbases/s and the compression ratio.
    python tools/cli_bench.py [--config C2] [--reads N] [--threads T]
"""
import argparse
import json
import os
import subprocess
import sys
import tempfile
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from ntcomp_b200 import synth  # noqa: E402

CLI = os.path.join(ROOT, "ntcomp_b200", "bin", "ntcomp-b200")


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--config", default="C2")
    ap.add_argument("--reads", type=int, default=None)
    ap.add_argument("--threads", type=int, default=0)
    ap.add_argument("--tmp", default=None, help="scratch directory (default: the system temp dir)")
    args = ap.parse_args()
    cfg = dict(synth.CONFIGS[args.config])
    if args.reads:
        cfg["n_reads"] = args.reads
    tmp = tempfile.mkdtemp(prefix="ntc_cli_", dir=args.tmp)
    g = synth.genome(cfg["genome"])
    with open(os.path.join(tmp, "ref.fa"), "wb") as f:
        f.write(b">genome\n" + bytes(g) + b"\n")
    bases, off = synth.reads(g, cfg["seed"], 0, cfg["n_reads"], cfg["read_len"], cfg["err_ppm"])
    L = cfg["read_len"]
    import numpy as np
    rows = np.empty((cfg["n_reads"], L + 1), dtype=np.uint8)
    rows[:, :L] = bases.reshape(-1, L)
    rows[:, L] = 10
    with open(os.path.join(tmp, "reads.fa"), "wb") as f:      # FASTA with constant headers (header text is not stored)
        hdr = np.frombuffer(b">r\n", dtype=np.uint8)
        out = np.empty((cfg["n_reads"], 3 + L + 1), dtype=np.uint8)
        out[:, :3] = hdr
        out[:, 3:] = rows
        f.write(out.tobytes())
    prefix = os.path.join(tmp, "idx")
    thr = ["-t", str(args.threads)] if args.threads else []
    t0 = time.time()
    subprocess.run([CLI, "build", "-k", str(cfg["k"]), "-o", prefix, os.path.join(tmp, "ref.fa")] + thr, check=True, capture_output=True)
    t1 = time.time()
    with open(os.path.join(tmp, "encoded.dat"), "wb") as f:
        r = subprocess.run([CLI, "encode", "-i", prefix, os.path.join(tmp, "reads.fa")] + thr, stdout=f, stderr=subprocess.PIPE)
    if r.returncode:
        raise SystemExit("encode failed: " + r.stderr.decode())
    enc_err = r.stderr.decode()
    sys.stderr.write(enc_err)
    enc_log = enc_err.strip().splitlines()[-1]
    t2 = time.time()
    with open(os.path.join(tmp, "decoded.fa"), "wb") as f:
        r = subprocess.run([CLI, "decode", "-i", prefix, os.path.join(tmp, "encoded.dat")] + thr, stdout=f, stderr=subprocess.PIPE)
    if r.returncode:
        raise SystemExit("decode failed: " + r.stderr.decode())
    dec_err = r.stderr.decode()
    sys.stderr.write(dec_err)
    dec_log = dec_err.strip().splitlines()[-1]
    t3 = time.time()
    want = b"".join(b">seq.%d\n%s\n" % (i + 1, bytes(bases[i * L:(i + 1) * L])) for i in range(min(cfg["n_reads"], 1000)))
    got = open(os.path.join(tmp, "decoded.fa"), "rb").read()
    n_bases = cfg["n_reads"] * L
    ok = got.startswith(want) and got.endswith(bytes(bases[-L:]) + b"\n") and got.count(b"\n") == 2 * cfg["n_reads"]
    import hashlib
    import re
    def pipe_seconds(log):               # "... | 0.41 s (+0.90 s index)"
        m = re.search(r"\| ([0-9.]+) s \(\+([0-9.]+) s index\)", log)
        return (float(m.group(1)), float(m.group(2))) if m else (None, None)
    def traces(err):                     # NTC_TRACE=1: the stage lines, without the per-call ones
        return [ln[len("[ntc trace] "):] for ln in err.splitlines()
                if ln.startswith("[ntc trace]") and "encode_columns ms" not in ln and "decode_columns_fasta ms" not in ln]
    enc_s, enc_setup = pipe_seconds(enc_log)
    dec_s, dec_setup = pipe_seconds(dec_log)
    line = {"config": args.config, "reads": cfg["n_reads"], "bases": n_bases, "round_trip_ok": ok,
            "encode_pipeline_s": enc_s, "encode_setup_s": enc_setup, "decode_pipeline_s": dec_s, "decode_setup_s": dec_setup,
            "encode_gbases_per_s": n_bases / enc_s / 1e9 if enc_s else None, "decode_gbases_per_s": n_bases / dec_s / 1e9 if dec_s else None,
            "encoded_sha256": hashlib.sha256(open(os.path.join(tmp, "encoded.dat"), "rb").read()).hexdigest(),
            "trace": {"encode": traces(enc_err), "decode": traces(dec_err)},
            "env": {k: v for k, v in os.environ.items() if k.startswith("NTC_")},
            "build_s": t1 - t0, "encode_wall_s": t2 - t1, "decode_wall_s": t3 - t2,
            "encode_log": enc_log, "decode_log": dec_log,
            "encoded_bytes": os.path.getsize(os.path.join(tmp, "encoded.dat")),
            "bits_per_base": 8.0 * os.path.getsize(os.path.join(tmp, "encoded.dat")) / n_bases,
            "host_threads": args.threads or os.cpu_count()}
    print(json.dumps(line))
    for fn in os.listdir(tmp):
        os.remove(os.path.join(tmp, fn))
    os.rmdir(tmp)


if __name__ == "__main__":
    main()
