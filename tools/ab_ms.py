"""A/B timing of the encode / decode kernels for one library build (development aid; bench.py is the contract).

  NTC_LIB_PATH=build/variant.so NTC_L2_PERSIST_MB=0 python tools/ab_ms.py [label] [n_reads] [read_len] [err_ppm] [G]

Prints one line per run with the per-kernel-class CUDA-event times of ctx.kernel_times() and a record-stream checksum
(so variants can be compared for identical output without the oracle)."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import ntcomp_b200 as nt
from ntcomp_b200 import synth, _lib

label = sys.argv[1] if len(sys.argv) > 1 else "run"
n_reads = int(sys.argv[2]) if len(sys.argv) > 2 else 10_000_000
L = int(sys.argv[3]) if len(sys.argv) > 3 else 150
err = int(sys.argv[4]) if len(sys.argv) > 4 else 10_000
G = int(sys.argv[5]) if len(sys.argv) > 5 else 5_000_000
k = 91
path = "/tmp/ab_index_%d_%d.ntci" % (G, k)
g = synth.genome(G)
if os.path.exists(path):
    idx = nt.Index.load(path)
else:
    idx = nt.Index.build([g], k)
    idx.save(path)
dev = torch.device("cuda:0")
d_g = torch.from_numpy(g).to(dev)
nb = n_reads * L
ctx = nt.Context(idx, 0)
st = torch.cuda.current_stream().cuda_stream
d_bases = torch.empty(nb, dtype=torch.uint8, device=dev)
_lib.check(_lib.lib().ntc_synth_reads_device(ctx._h, d_g.data_ptr(), G, 999, 0, n_reads, L, err, d_bases.data_ptr(), st))
d_off = torch.arange(n_reads + 1, dtype=torch.int64, device=dev) * L
cap = nb // 4
d_rec = torch.empty(cap, dtype=torch.int64, device=dev)
d_roff = torch.empty(n_reads + 1, dtype=torch.int64, device=dev)
d_cnt = torch.zeros(4, dtype=torch.int64, device=dev)
d_out = torch.empty(nb + n_reads * 20 + 64, dtype=torch.uint8, device=dev)


def enc():
    ctx.encode_device(d_bases.data_ptr(), d_off.data_ptr(), n_reads, nb, d_rec.data_ptr(), cap, d_roff.data_ptr(), d_cnt.data_ptr(), st)


def dec(n_rec):
    ctx.decode_device(d_rec.data_ptr(), n_rec, 1, True, d_out.data_ptr(), d_out.numel(), 0, 0, d_cnt.data_ptr(), st)


ctx.set_profiling(True)
for _ in range(3):
    enc()
ctx.status(st)
n_rec = int(d_cnt[0].item())
for _ in range(3):
    dec(n_rec)
ctx.status(st)
ctx.kernel_times(reset=True)
reps = 5
for _ in range(reps):
    enc()
    dec(n_rec)
ctx.status(st)
t = ctx.kernel_times()
chk = int((d_rec[:n_rec] * 0x9E3779B97F4A7C15).sum().item()) & 0xFFFFFFFFFFFF
nbytes = int(d_cnt[1].item()) // 8 * 8
wv = d_out[:nbytes].view(torch.int64)
ochk = int((wv * (torch.arange(wv.numel(), device=dev) % 1021 + 1)).sum().item()) & 0xFFFFFFFFFFFF
parts = " ".join("%s=%.3f" % (n, v[0] / reps) for n, v in t.items() if v[1])
tot = sum(v[0] for v in t.values()) / reps
print("%-14s reads=%d L=%d err=%d records=%d chk=%012x out=%012x  total=%.3f ms  %s" % (label, n_reads, L, err, n_rec, chk, ochk, tot, parts), flush=True)
