#!/usr/bin/env python
"""Generate the compare-exchange networks used by the column-select kernel (K1).

Writes csrc/gen/networks_<part>.cu: straight-line sorting / selection networks over a
register array ``float v[N]``, one explicit template
specialisation of ``flagg::Net<N>`` per size, spread over several translation units so
``make -j`` compiles them in parallel.

Two networks per size N:
  * ``sort``    a full sorting network (Batcher odd-even merge sort on the next power
                of two, comparators touching the +inf padding removed) -- used for any
                runtime n <= N, any rank, and for the trimmed mean;
  * ``median``  the same network pruned backwards from output wire (N-1)//2, with
                half comparators (only min or only max kept) where one output is
                dead -- used when n == N exactly.

Sizes: every N <= 32, then multiples of 4 up to 64, multiples of 8 up to 128, plus the
headline n = 100 (BASELINE.json config 3).  The list is mirrored in colselect.cu via the
generated header gen/network_sizes.h.
"""
import os
import sys

SIZES = sorted(set(list(range(2, 33)) + list(range(36, 65, 4)) + list(range(72, 129, 8)) + [100]))
PARTS = 8


def oddeven_merge_sort(n):
    p = 1
    while p < n:
        p *= 2
    ces = []

    def merge(lo, hi, r):
        step = r * 2
        if step < hi - lo:
            merge(lo, hi, step)
            merge(lo + r, hi, step)
            for i in range(lo + r, hi - r, step):
                ces.append((i, i + r))
        else:
            ces.append((lo, lo + r))

    def sort(lo, hi):
        if hi - lo >= 1:
            mid = lo + (hi - lo) // 2
            sort(lo, mid)
            sort(mid + 1, hi)
            merge(lo, hi, 1)

    sort(0, p - 1)
    return [(a, b) for a, b in ces if a < n and b < n]


def prune_to_outputs(ces, outs):
    """Backward liveness: returns [(a, b, keep_lo, keep_hi)] that still determine `outs`."""
    live = set(outs)
    kept = []
    for a, b in reversed(ces):
        lo, hi = a in live, b in live
        if lo or hi:
            kept.append((a, b, lo, hi))
            live.add(a)
            live.add(b)
    return kept[::-1]


def verify(n, ops, outs, trials=200):
    import random
    rnd = random.Random(n)
    for _ in range(trials):
        v = [rnd.randint(0, max(1, n // 2)) for _ in range(n)]
        s = sorted(v)
        w = v[:]
        for a, b, lo, hi in ops:
            x, y = min(w[a], w[b]), max(w[a], w[b])
            if lo:
                w[a] = x
            if hi:
                w[b] = y
        for o in outs:
            assert w[o] == s[o], (n, o)


def mix_for(n):
    """(num, den): share of full compare-exchanges whose max goes to the FMA pipe.  Tuned on B200
    (tools/k1_tune): large networks are ALU/issue co-bound and want ~3/5; small ones are HBM-bound."""
    env = os.environ.get("FLAGG_MIX")
    if env:
        a, b = env.split("/")
        return int(a), int(b)
    return (2, 3) if n >= 48 else (5, 8)


def emit(n):
    full = oddeven_merge_sort(n)
    full_ops = [(a, b, True, True) for a, b in full]
    verify(n, full_ops, range(n))
    r = (n - 1) // 2
    med = prune_to_outputs(full, [r])
    verify(n, med, [r])

    def body(ops):
        # A fraction MIX of the full compare-exchanges computes its upper output on the FMA pipe
        # (CE1, two IMADs) instead of the ALU pipe (CE0, FMNMX); spread evenly (Bresenham).
        out = []
        nfull = 0
        num, den = mix_for(n)
        for a, b, lo, hi in ops:
            if lo and hi:
                on_fma = ((nfull + 1) * num) // den - (nfull * num) // den
                out.append(f"CE{on_fma}({a},{b})")
                nfull += 1
            else:
                out.append(f"{'CE_LO' if lo else 'CE_HI'}({a},{b})")
        lines = []
        for i in range(0, len(out), 8):
            lines.append("    " + " ".join(out[i:i + 8]))
        return "\n".join(lines)

    # runtime n in (prev, N] (prev = next smaller generated size): rank (n-1)//2 lies in [prev//2, (N-1)//2]
    prev = max([m for m in SIZES if m < n], default=1)
    rmin = prev // 2
    med_any = prune_to_outputs(full, list(range(rmin, r + 1)))
    verify(n, med_any, range(rmin, r + 1))
    nops_med = sum(int(lo) + int(hi) for _, _, lo, hi in med)
    txt = f"// N={n}: sort {len(full)} CE ({2 * len(full)} min/max); median(rank {r}) {len(med)} CE ({nops_med} min/max)\n"
    txt += f"template <> __device__ __forceinline__ void Net<{n}>::sort(float (&v)[{n}], int one, int mone) {{\n{body(full_ops)}\n}}\n"
    txt += f"template <> __device__ __forceinline__ float Net<{n}>::median(float (&v)[{n}], int one, int mone) {{\n{body(med)}\n    return v[{r}];\n}}\n"
    txt += f"template <> __device__ __forceinline__ void Net<{n}>::median_any(float (&v)[{n}], int one, int mone) {{\n{body(med_any)}\n}}\n"
    return txt, len(full), len(med), nops_med


def main():
    global SIZES
    here = os.path.dirname(os.path.abspath(__file__))
    gen = os.path.join(here, "gen")
    if "--out" in sys.argv:
        gen = os.path.join(here, sys.argv[sys.argv.index("--out") + 1])
    if "--sizes" in sys.argv:  # tuning builds: only these network sizes
        SIZES = sorted(int(s) for s in sys.argv[sys.argv.index("--sizes") + 1].split(","))
    os.makedirs(gen, exist_ok=True)
    parts = [[] for _ in range(PARTS)]
    # balance by approximate cost n*log2(n)^2, largest first
    import math
    load = [0.0] * PARTS
    for n in sorted(SIZES, key=lambda s: -s):
        i = load.index(min(load))
        parts[i].append(n)
        load[i] += n * max(1.0, math.log2(n)) ** 2
    stats = []
    for i, sizes in enumerate(parts):
        with open(os.path.join(gen, f"networks_{i}.cu"), "w") as fh:
            fh.write("// GENERATED by gen_networks.py -- do not edit.\n#include \"colselect_net.cuh\"\nnamespace flagg {\n")
            for n in sorted(sizes):
                txt, nf, nm, nom = emit(n)
                stats.append((n, nf, nm, nom))
                fh.write(txt)
            fh.write("}  // namespace flagg\n")
            for n in sorted(sizes):
                fh.write(f"FLAGG_INSTANTIATE_COLSELECT({n})\n")
    with open(os.path.join(gen, "network_sizes.h"), "w") as fh:
        fh.write("// GENERATED by gen_networks.py -- do not edit.\n#pragma once\n")
        fh.write("#define FLAGG_NET_SIZES(X) " + " ".join(f"X({n})" for n in SIZES) + "\n")
        fh.write(f"#define FLAGG_NET_MAX {max(SIZES)}\n")
        # next smaller generated size: a network of size N serves n in (net_prev(N), N]
        chain = " ".join(f"n == {n} ? {max([m for m in SIZES if m < n], default=1)} :" for n in SIZES)
        fh.write(f"__host__ __device__ constexpr int net_prev(int n) {{ return {chain} 1; }}\n")
    if "-v" in sys.argv:
        for n, nf, nm, nom in sorted(stats):
            print(f"N={n:4d} sort={nf:5d} CE  median={nm:5d} CE / {nom:5d} min-max ops")
    print(f"generated {len(SIZES)} network sizes in {PARTS} files under {gen}")


if __name__ == "__main__":
    main()
