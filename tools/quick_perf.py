"""Quick device-resident timing of the encode/decode pipelines (development aid; bench.py is the contract)."""
import sys, time, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import ntcomp_b200 as nt
from ntcomp_b200 import synth

G = int(sys.argv[1]) if len(sys.argv) > 1 else 5_000_000
k = int(sys.argv[2]) if len(sys.argv) > 2 else 31
n_reads = int(sys.argv[3]) if len(sys.argv) > 3 else 1_000_000
read_len = int(sys.argv[4]) if len(sys.argv) > 4 else 150
err = int(sys.argv[5]) if len(sys.argv) > 5 else 10_000

t = time.time(); g = synth.genome(G); idx = nt.Index.build([g], k); print("index build %.2fs n=%d" % (time.time() - t, idx.n_sets))
t = time.time(); ctx = nt.Context(idx, 0); print("ctx create %.2fs" % (time.time() - t), ctx.index_bytes())
dev = torch.device("cuda:0")
d_g = torch.from_numpy(g).to(dev)
n_bases = n_reads * read_len
d_bases = torch.empty(n_bases, dtype=torch.uint8, device=dev)
from ntcomp_b200 import _lib
_lib.check(_lib.lib().ntc_synth_reads_device(ctx._h, d_g.data_ptr(), G, 999, 0, n_reads, read_len, err, d_bases.data_ptr(), 0))
d_off = (torch.arange(n_reads + 1, dtype=torch.int64, device=dev) * read_len)
cap = n_bases // 4
d_rec = torch.empty(cap, dtype=torch.int64, device=dev)
d_roff = torch.empty(n_reads + 1, dtype=torch.int64, device=dev)
d_cnt = torch.zeros(4, dtype=torch.int64, device=dev)
d_out = torch.empty(n_bases + n_reads * 20 + 64, dtype=torch.uint8, device=dev)
st = torch.cuda.current_stream().cuda_stream
def enc():
    ctx.encode_device(d_bases.data_ptr(), d_off.data_ptr(), n_reads, n_bases, d_rec.data_ptr(), cap, d_roff.data_ptr(), d_cnt.data_ptr(), st)
def dec(n_rec, fasta):
    ctx.decode_device(d_rec.data_ptr(), n_rec, 1, fasta, d_out.data_ptr(), d_out.numel(), 0, 0, d_cnt.data_ptr(), st)
enc(); ctx.status(st); n_rec = int(d_cnt[0].item()); print("records", n_rec, "per read %.2f" % (n_rec / n_reads))
for name, fn in [("encode", enc), ("decode", lambda: dec(n_rec, False)), ("decode_fasta", lambda: dec(n_rec, True))]:
    for _ in range(2): fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    reps = 5
    e0.record()
    for _ in range(reps): fn()
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / reps
    ctx.status(st)
    print("%s: %.3f ms  %.1f Mbases/s" % (name, ms, n_bases / ms / 1e3))
dec(n_rec, False); ctx.status(st)
ok = bool((d_out[:n_bases] == d_bases).all().item())
print("roundtrip", ok)
