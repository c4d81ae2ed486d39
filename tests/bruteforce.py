"""Definition-level SBWT model in plain Python strings (tiny inputs only).

Independent of oracle/*.cpp: node set, colex order, subset rows, C, LCS and matching statistics are
computed from their textbook definitions with string operations, so that the C++ oracle's bit-packed
construction can be cross-checked.  (This is a second restatement, not the Rust crate: parity with
sbwt 0.3.11 stays unpinned, see oracle/oracle_index.h.)
"""

COMP = {"A": "T", "C": "G", "G": "C", "T": "A"}


def revcomp(s):
    return "".join(COMP[c] for c in reversed(s))


def split_acgt(seq):
    out, cur = [], []
    for ch in seq:
        if ch in "ACGT":
            cur.append(ch)
        else:
            if cur:
                out.append("".join(cur))
            cur = []
    if cur:
        out.append("".join(cur))
    return out


class BruteSBWT:
    def __init__(self, seqs, k, add_revcomp=True):
        self.k = k
        kmers = set()
        for seq in seqs:
            for piece in split_acgt(seq):
                strands = [piece, revcomp(piece)] if add_revcomp else [piece]
                for s in strands:
                    for i in range(len(s) - k + 1):
                        kmers.add(s[i:i + k])
        suffixes = {x[1:] for x in kmers}
        nodes = set(kmers)
        nodes.add("$" * k)
        for x in kmers:
            if x[:-1] not in suffixes:
                for j in range(1, k):
                    nodes.add("$" * j + x[:k - j])
        self.nodes = sorted(nodes, key=lambda s: s[::-1])  # '$' < 'A' < 'C' < 'G' < 'T' in ASCII
        self.n = len(self.nodes)
        node_set = set(self.nodes)
        self.rows = [[0] * self.n for _ in range(4)]
        seen_suffix = set()
        for i, u in enumerate(self.nodes):
            suf = u[1:]
            if suf in seen_suffix:
                continue  # only the first node of a (k-1)-suffix group carries labels
            seen_suffix.add(suf)
            for c, ch in enumerate("ACGT"):
                if suf + ch in node_set:
                    self.rows[c][i] = 1
        self.C = [1]
        for c in range(4):
            self.C.append(self.C[-1] + sum(self.rows[c]))
        self.lcs = [0] * self.n
        for i in range(1, self.n):
            a, b = self.nodes[i - 1], self.nodes[i]
            m = 0
            while m < k and a[k - 1 - m] == b[k - 1 - m] and a[k - 1 - m] != "$":
                m += 1
            self.lcs[i] = m

    def interval_of_suffix(self, suf):
        idx = [i for i, u in enumerate(self.nodes) if u.endswith(suf)]
        if not idx:
            return None
        assert idx == list(range(idx[0], idx[-1] + 1))
        return idx[0], idx[-1] + 1

    def matching_statistics(self, read):
        out = []
        for j in range(len(read)):
            best = (0, (0, self.n))
            for d in range(min(self.k, j + 1), 0, -1):
                iv = self.interval_of_suffix(read[j + 1 - d:j + 1])
                if iv is not None:
                    best = (d, iv)
                    break
            out.append(best)
        return out
