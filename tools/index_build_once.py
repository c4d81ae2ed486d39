"""one ntc_index_build_gpu call on a BASELINE genome (for ncu launch lists / captures); prints the phase stats"""
import json
import sys

sys.path.insert(0, __file__.rsplit("/", 2)[0])
import ntcomp_b200 as nt  # noqa: E402
from ntcomp_b200 import synth  # noqa: E402

G = int(sys.argv[1]) if len(sys.argv) > 1 else 5_000_000
k = int(sys.argv[2]) if len(sys.argv) > 2 else 91
st = {}
idx = nt.Index.build_gpu([synth.genome(G)], k, stats=st)
st["n_sets"] = idx.n_sets
print(json.dumps(st))
