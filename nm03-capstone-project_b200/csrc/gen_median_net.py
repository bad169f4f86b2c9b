#!/usr/bin/env python3
"""Generator for the min/max selection networks used by the fused preprocess kernel (K1).

The median window is evaluated as a data-oblivious network of min/max operations on packed
int16x2 registers (two pixels per lane, VIMNMX.S16x2 / VIMNMX3.S16x2 on sm_100a).  Networks are built
here as SSA programs, dead-code-eliminated from the one output rank that is needed, checked
against numpy on random inputs with heavy ties, and emitted as straight-line CUDA.

Building blocks:
  sort_n         — optimal small sorting networks (column presort)
  oe_merge       — Batcher odd-even merge of two sorted lists of any lengths (Knuth 5.3.4)
  kth_of_two     — k-th smallest of two sorted lists, min over i+j=k of max(A_i, B_j)
  pruning        — an element of a partial merge of m out of n elements can only be the final
                   rank-r element if its partial rank lies in [r-(n-m), r]; others are dropped.

Usage:  gen_median_net.py --emit > median_net.inc     (also: --report for op counts)
"""
import argparse
import itertools
import sys

import numpy as np


class Net:
    """SSA min/max program."""

    def __init__(self):
        self.ops = []      # (op, a, b[, c]) ; value id = n_inputs + index
        self.n_in = 0
        self.cache = {}

    def inputs(self, n):
        ids = list(range(self.n_in, self.n_in + n))
        assert not self.ops, "declare inputs first"
        self.n_in += n
        return ids

    def _emit(self, key):
        if key in self.cache:
            return self.cache[key]
        self.ops.append(key)
        vid = self.n_in + len(self.ops) - 1
        self.cache[key] = vid
        return vid

    def mn(self, a, b):
        if a == b:
            return a
        return self._emit(("min",) + tuple(sorted((a, b))))

    def mx(self, a, b):
        if a == b:
            return a
        return self._emit(("max",) + tuple(sorted((a, b))))

    def ce(self, a, b):
        return self.mn(a, b), self.mx(a, b)

    # ---- analysis ----
    def live_ops(self, outputs):
        live = set()
        stack = [o for o in outputs if o >= self.n_in]
        while stack:
            v = stack.pop()
            if v in live:
                continue
            live.add(v)
            for s in self.ops[v - self.n_in][1:]:
                if s >= self.n_in:
                    stack.append(s)
        return live

    def count(self, outputs):
        return len(self.live_ops(outputs))

    def simulate(self, in_arrays, outputs):
        vals = {i: a for i, a in enumerate(in_arrays)}
        live = self.live_ops(outputs)
        for idx, op in enumerate(self.ops):
            vid = self.n_in + idx
            if vid not in live:
                continue
            f = np.minimum if op[0].startswith("min") else np.maximum
            r = f(vals[op[1]], vals[op[2]])
            if len(op) == 4:
                r = f(r, vals[op[3]])
            vals[vid] = r
        return [vals[o] for o in outputs]

    def fuse3(self, outputs):
        """Fold min(min(a,b),c) -> min3(a,b,c) where the inner value has a single use (VIMNMX3)."""
        live = self.live_ops(outputs)
        uses = {}
        for v in live:
            for s in self.ops[v - self.n_in][1:]:
                uses[s] = uses.get(s, 0) + 1
        for o in outputs:
            uses[o] = uses.get(o, 0) + 1
        new_ops = list(self.ops)
        dead = set()
        for v in sorted(live):
            op = new_ops[v - self.n_in]
            if len(op) != 3:
                continue
            for pos in (1, 2):
                s = op[pos]
                if s >= self.n_in and s not in dead and uses.get(s, 0) == 1:
                    sop = new_ops[s - self.n_in]
                    if len(sop) == 3 and sop[0] == op[0]:
                        other = op[2] if pos == 1 else op[1]
                        new_ops[v - self.n_in] = (op[0] + "3", sop[1], sop[2], other)
                        dead.add(s)
                        break
        self.ops = new_ops
        return self

    def emit_cuda(self, outputs, in_names, out_names, indent="  ", prefix="t", order="build"):
        live = self.live_ops(outputs)
        name = {i: n for i, n in enumerate(in_names)}
        lines = []
        fn = {"min": "NM03_MIN", "max": "NM03_MAX", "min3": "NM03_MIN3", "max3": "NM03_MAX3"}
        seq = list(enumerate(self.ops))
        if order == "level":   # as-soon-as-possible levels: maximal instruction-level parallelism in source order
            depth = {}
            for idx, op in seq:
                depth[self.n_in + idx] = 1 + max([depth.get(a, 0) for a in op[1:]])
            seq.sort(key=lambda t: (depth[self.n_in + t[0]], t[0]))
        for idx, op in seq:
            vid = self.n_in + idx
            if vid not in live:
                continue
            name[vid] = f"{prefix}{idx}"
            args = ", ".join(name[s] for s in op[1:])
            lines.append(f"{indent}const uint32_t {name[vid]} = {fn[op[0]]}({args});")
        for o, on in zip(outputs, out_names):
            lines.append(f"{indent}{on} = {name[o]};")
        return "\n".join(lines)


# optimal sorting networks (comparator lists) for n <= 8
SORT_NETS = {
    1: [],
    2: [(0, 1)],
    3: [(0, 2), (0, 1), (1, 2)],
    4: [(0, 2), (1, 3), (0, 1), (2, 3), (1, 2)],
    5: [(0, 3), (1, 4), (0, 2), (1, 3), (0, 1), (2, 4), (1, 2), (3, 4), (2, 3)],
    6: [(0, 5), (1, 3), (2, 4), (1, 2), (3, 4), (0, 3), (2, 5), (0, 1), (2, 3), (4, 5), (1, 2), (3, 4)],
    7: [(0, 6), (2, 3), (4, 5), (0, 2), (1, 4), (3, 6), (0, 1), (2, 5), (3, 4), (1, 2), (4, 6), (2, 3), (4, 5),
        (1, 2), (3, 4), (5, 6)],
}


def sort_n(net, vals):
    v = list(vals)
    if len(v) in SORT_NETS:
        for i, j in SORT_NETS[len(v)]:
            v[i], v[j] = net.ce(v[i], v[j])
        return v
    half = len(v) // 2
    return oe_merge(net, sort_n(net, v[:half]), sort_n(net, v[half:]))


def oe_merge(net, A, B):
    """Batcher odd-even merge, arbitrary lengths (Knuth TAOCP 5.3.4)."""
    if not A:
        return list(B)
    if not B:
        return list(A)
    if len(A) == 1 and len(B) == 1:
        return list(net.ce(A[0], B[0]))
    v = oe_merge(net, A[0::2], B[0::2])
    w = oe_merge(net, A[1::2], B[1::2])
    out = [v[0]]
    for i in range(len(w)):
        if i + 1 < len(v):
            lo, hi = net.ce(w[i], v[i + 1])
            out += [lo, hi]
        else:
            out.append(w[i])
    if len(v) > len(w) + 1:
        out += v[len(w) + 1:]
    return out


def kth_of_two(net, A, B, k):
    """k-th smallest (1-based) of the union of sorted A and B: min over i+j=k of max(A_i, B_j)."""
    terms = []
    for i in range(0, k + 1):
        j = k - i
        if i > len(A) or j > len(B):
            continue
        if i == 0:
            terms.append(B[j - 1])
        elif j == 0:
            terms.append(A[i - 1])
        else:
            terms.append(net.mx(A[i - 1], B[j - 1]))
    while len(terms) > 1:  # balanced min tree
        nxt = [net.mn(terms[i], terms[i + 1]) for i in range(0, len(terms) - 1, 2)]
        if len(terms) & 1:
            nxt.append(terms[-1])
        terms = nxt
    return terms[0]


class Part:
    """A sorted, pruned partial merge: `vals` are the elements whose rank within the part is
    lo+1 .. lo+len(vals) (1-based); `size` elements went in."""

    def __init__(self, vals, size, lo=0):
        self.vals, self.size, self.lo = vals, size, lo


def prune(part, total, rank):
    """Keep only the elements that can still be the rank-th (1-based) of `total`."""
    rest = total - part.size
    keep_lo = max(1, rank - rest)       # partial rank must be >= rank - rest
    keep_hi = min(part.size, rank)      # and <= rank
    vals = []
    for idx, v in enumerate(part.vals):
        r = part.lo + idx + 1
        if keep_lo <= r <= keep_hi:
            vals.append(v)
    assert len(vals) == keep_hi - keep_lo + 1, "pruned range left the kept window"
    return Part(vals, part.size, keep_lo - 1)


def merge_parts(net, P, Q, total, rank):
    """Merge two pruned parts.  Correctness: a dropped element can never sit at position `rank` of the
    final order, so it may be replaced by -inf (dropped low) or +inf (dropped high) without changing
    the answer; in that modified multiset the merged kept list holds exactly the consecutive ranks
    P.lo + Q.lo + 1 ... and can be pruned again by the same rule."""
    merged = oe_merge(net, P.vals, Q.vals)
    return prune(Part(merged, P.size + Q.size, P.lo + Q.lo), total, rank)


def select_parts(net, P, Q, total, rank):
    """Final step: P and Q together hold `total` elements; return the rank-th."""
    assert P.size + Q.size == total
    k = rank - P.lo - Q.lo
    return kth_of_two(net, P.vals, Q.vals, k)


# ---------------------------------------------------------------------------------------------
# designs
# ---------------------------------------------------------------------------------------------
def dyadic_blocks(lo, hi, off, maxlen):
    """Decompose the column range [lo, hi) into blocks of power-of-two length aligned at `off`."""
    blocks = []
    x = lo
    while x < hi:
        ln = 1
        while ln * 2 <= maxlen and (x - off) % (ln * 2) == 0 and x + ln * 2 <= hi:
            ln *= 2
        blocks.append((x, ln))
        x += ln
    return blocks


def run_design(R, k, off, maxlen, presort=True):
    """R horizontally adjacent outputs computed by one thread.

    Inputs: (R+k-1) columns of k raw samples (column c = window columns of output i for c in
    [i, i+k)).  Each column is sorted once (shared by up to k outputs), aligned groups of 2/4 columns
    are merged once and shared by the outputs that contain them, every partial merge is pruned to
    the ranks that can still be the median, and the last step is a direct rank selection."""
    net = Net()
    ncol = R + k - 1
    raw = net.inputs(ncol * k)
    total, rank = k * k, (k * k + 1) // 2
    cols = []
    for c in range(ncol):
        cols += sort_n(net, raw[c * k:(c + 1) * k]) if presort else raw[c * k:(c + 1) * k]
    cache = {}

    def block(x, ln):
        if (x, ln) not in cache:
            if ln == 1:
                cache[(x, ln)] = prune(Part(cols[x * k:(x + 1) * k], k), total, rank)
            else:
                cache[(x, ln)] = merge_parts(net, block(x, ln // 2), block(x + ln // 2, ln // 2), total, rank)
        return cache[(x, ln)]

    outs = []
    for i in range(R):
        parts = [block(x, ln) for x, ln in dyadic_blocks(i, i + k, off, maxlen)]
        while len(parts) > 2:
            parts.sort(key=lambda p: p.size)
            a, b = parts[0], parts[1]
            parts = parts[2:] + [merge_parts(net, a, b, total, rank)]
        if len(parts) == 2:
            outs.append(select_parts(net, parts[0], parts[1], total, rank))
        else:
            outs.append(parts[0].vals[rank - 1 - parts[0].lo])
    return net, raw, outs


def run_design_core(R, k, pair_off, blk_off, maxlen, presort=True):
    """Design B: outputs are handled in adjacent pairs (i, i+1) that share a CORE of k-1 columns
    [i+1, i+k).  The core is merged once from aligned blocks of columns (shared with the neighbouring
    pairs) and pruned to the few ranks that can still be the median (8 of 42 for k=7); each output
    then needs only one rank selection between the core and its one private column."""
    net = Net()
    ncol = R + k - 1
    raw = net.inputs(ncol * k)
    total, rank = k * k, (k * k + 1) // 2
    cols = []
    for c in range(ncol):
        cols += sort_n(net, raw[c * k:(c + 1) * k]) if presort else raw[c * k:(c + 1) * k]
    cache = {}

    def block(x, ln):
        if (x, ln) not in cache:
            if ln == 1:
                cache[(x, ln)] = prune(Part(cols[x * k:(x + 1) * k], k), total, rank)
            else:
                cache[(x, ln)] = merge_parts(net, block(x, ln // 2), block(x + ln // 2, ln // 2), total, rank)
        return cache[(x, ln)]

    cores = {}

    def core(a, b):
        if (a, b) not in cores:
            parts = [block(x, ln) for x, ln in dyadic_blocks(a, b, blk_off, maxlen)]
            while len(parts) > 1:
                parts.sort(key=lambda p: p.size)
                parts = parts[2:] + [merge_parts(net, parts[0], parts[1], total, rank)]
            cores[(a, b)] = parts[0]
        return cores[(a, b)]

    outs = []
    for i in range(R):
        if (i - pair_off) % 2 == 0 and i + 1 < R:
            c, single = core(i + 1, i + k), i            # first of a pair
        elif (i - pair_off) % 2 == 1 and i - 1 >= 0:
            c, single = core(i, i + k - 1), i + k - 1    # second of a pair
        else:                                            # unpaired output at the edge of the run
            c, single = core(i + 1, i + k), i
        outs.append(select_parts(net, c, block(single, 1), total, rank))
    return net, raw, outs


def best_core_design(R, k):
    best = None
    for maxlen in (1, 2, 4, 8):
        for blk_off in range(maxlen):
            for pair_off in (0, 1):
                net, raw, outs = run_design_core(R, k, pair_off, blk_off, maxlen)
                n = net.count(outs)
                if best is None or n < best[0]:
                    best = (n, pair_off, blk_off, maxlen)
    return best


def best_run_design(R, k):
    best = None
    for maxlen in (1, 2, 4, 8):
        for off in range(maxlen):
            net, raw, outs = run_design(R, k, off, maxlen)
            n = net.count(outs)
            if best is None or n < best[0]:
                best = (n, maxlen, off)
    return best


def check_run(net, raw, outs, R, k, trials=3000):
    rng = np.random.default_rng(1)
    ncol = R + k - 1
    for hi in (3, 60, 32000):
        cols = rng.integers(-hi, hi, (ncol, k, trials)).astype(np.int32)
        arrays = [cols[c, r] for c in range(ncol) for r in range(k)]
        got = net.simulate(arrays, outs)
        for i in range(R):
            exp = np.sort(cols[i:i + k].reshape(k * k, trials), axis=0)[(k * k) // 2]
            if not np.array_equal(got[i], exp):
                return False
    return True


RUN = 8  # outputs per thread-run (packed registers: 16 pixels)
ORDER = "build"


RUN_FOR = {3: 8, 5: 8, 7: 8, 9: 2}  # outputs per thread-run for each window (9x9 is limited by registers)


def emit(args):
    print("// GENERATED by gen_median_net.py --emit — do not edit.")
    print("// Straight-line min/max networks on packed int16x2 registers (two pixels per lane).")
    print("// NM03_MEDIAN_RUN_<k>(v, o): v[c][r] = raw sample r of column c, c < RUNLEN+k-1, r < k;")
    print("// o[i], i < RUNLEN = median of the k x k window made of columns i .. i+k-1 (RUNLEN = NM03_MEDIAN_RUNLEN_<k>).")
    for k in (3, 5, 7, 9):
        run = RUN if (args.run_given and k != 9) else RUN_FOR[k]
        n, pair_off, blk_off, maxlen = best_core_design(run, k)
        net, raw, outs = run_design_core(run, k, pair_off, blk_off, maxlen)
        assert check_run(net, raw, outs, run, k, trials=3000 if k < 9 else 600), f"k={k} failed verification"
        net.fuse3(outs)
        assert check_run(net, raw, outs, run, k, trials=3000 if k < 9 else 600), f"k={k} failed verification after 3-input fusion"
        n3 = net.count(outs)
        print(f"\n// k={k}: {n} two-input ops, {n3} instructions after min3/max3 fusion "
              f"({n3 / (2.0 * run):.1f} per pixel); shared-core design, pair offset {pair_off}, "
              f"aligned blocks <= {maxlen} columns at offset {blk_off}")
        print(f"#define NM03_MEDIAN_RUNLEN_{k} {run}")
        print(f"#define NM03_MEDIAN_OPS_{k} {n3}")
        print(f"#define NM03_MEDIAN_RUN_{k}(v, o) do {{ \\")
        in_names = [f"(v)[{c}][{r}]" for c in range(run + k - 1) for r in range(k)]
        body = net.emit_cuda(outs, in_names, [f"(o)[{i}]" for i in range(run)], indent="  ", prefix=f"m{k}_", order=ORDER)
        for ln in body.split("\n"):
            print(ln + " \\")
        print("} while (0)")


def report(args):
    for k in (3, 5, 7):
        for R in (2, 4, 8, 16):
            n, pair_off, blk_off, maxlen = best_core_design(R, k)
            net, raw, outs = run_design_core(R, k, pair_off, blk_off, maxlen)
            ok = check_run(net, raw, outs, R, k, trials=500)
            net.fuse3(outs)
            print(f"core design k={k} R={R}: {n} ops ({n / R:.1f}/output), {net.count(outs)} fused "
                  f"({net.count(outs) / (2.0 * R):.1f} instr/pixel packed) pair_off={pair_off} "
                  f"blocks<={maxlen}@{blk_off} verified={ok}")
    for k in (3, 5, 7):
        for R in (1, 2, 4, 8):
            n, maxlen, off = best_run_design(R, k)
            net, raw, outs = run_design(R, k, off, maxlen)
            ok = check_run(net, raw, outs, R, k, trials=500)
            net.fuse3(outs)
            print(f"k={k} R={R}: {n} ops ({n / R:.1f}/output), {net.count(outs)} fused "
                  f"({net.count(outs) / (2.0 * R):.1f} instr/pixel packed) groups<={maxlen}@{off} verified={ok}")


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--emit", action="store_true")
    ap.add_argument("--report", action="store_true")
    ap.add_argument("--run", type=int, default=None, help="outputs per thread-run for windows 3/5/7 (tuning)")
    ap.add_argument("--order", default="build", choices=["build", "level"], help="emission order of the operations")
    a = ap.parse_args()
    a.run_given = a.run is not None
    if a.run_given:
        RUN = a.run
    ORDER = a.order
    if a.emit:
        emit(a)
    else:
        report(a)
