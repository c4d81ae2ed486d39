"""Sweeps ms_kernel's speculation window (NTC_SPEC_LO / NTC_SPEC_SPAN) on a C3-shaped batch (development aid)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import ntcomp_b200 as nt
from ntcomp_b200 import synth, _lib
G, k, n_reads, L = 5_000_000, int(sys.argv[1]) if len(sys.argv) > 1 else 91, 4_000_000, 150
g = synth.genome(G); idx = nt.Index.build([g], k)
dev = torch.device("cuda:0"); d_g = torch.from_numpy(g).to(dev)
nb = n_reads * L
d_bases = torch.empty(nb, dtype=torch.uint8, device=dev)
d_off = torch.arange(n_reads + 1, dtype=torch.int64, device=dev) * L
d_rec = torch.empty(nb // 6, dtype=torch.int64, device=dev); d_roff = torch.empty(n_reads + 1, dtype=torch.int64, device=dev)
d_cnt = torch.zeros(4, dtype=torch.int64, device=dev)
st = torch.cuda.current_stream().cuda_stream
first = True
for lo, span, od in [(10, 6, 0), (0, 0, 0), (11, 3, 0), (10, 3, 0), (11, 2, 0), (10, 4, 0), (11, 5, 0), (9, 7, 0), (12, 3, 0), (0, 255, 0)]:
    os.environ["NTC_SPEC_LO"] = str(lo); os.environ["NTC_SPEC_SPAN"] = str(span); os.environ["NTC_SPEC_ONDEMAND"] = str(od)
    ctx = nt.Context(idx, 0)
    if first:
        _lib.check(_lib.lib().ntc_synth_reads_device(ctx._h, d_g.data_ptr(), G, 999, 0, n_reads, L, 10_000, d_bases.data_ptr(), st)); first = False
    ctx.set_profiling(True)
    for _ in range(5):
        ctx.encode_device(d_bases.data_ptr(), d_off.data_ptr(), n_reads, nb, d_rec.data_ptr(), d_rec.numel(), d_roff.data_ptr(), d_cnt.data_ptr(), st)
    ctx.status(st); ctx.kernel_times(reset=True)
    for _ in range(5):
        ctx.encode_device(d_bases.data_ptr(), d_off.data_ptr(), n_reads, nb, d_rec.data_ptr(), d_rec.numel(), d_roff.data_ptr(), d_cnt.data_ptr(), st)
    ctx.status(st); t = ctx.kernel_times()
    print("ondemand=%d " % od, end=""); print("spec [%d,%d): ms_kernel %.3f ms  (%.1f Gbases/s)  records %d" % (lo, lo + span, t["ms"][0] / t["ms"][1], nb / (t["ms"][0] / t["ms"][1]) / 1e6, int(d_cnt[0].item())), flush=True)
    ctx.close()
