#!/usr/bin/env python
"""Aggregate the SASS page of an ncu report per handler of the VM interpreter kernel.

    ncu -i X.ncu-rep --page source --csv --print-source sass > x.csv
    python profiles/sass_regions.py x.csv [kernel-index]

Regions are the instruction runs between indirect jumps (BRX) - one per group of handlers - plus the
dispatch block, recognised as the instructions executed once per micro-op.  For each region: stall samples,
instructions executed, shared-memory wavefronts (and the excess over the ideal), the leading stall reasons
and the instruction mix.  Also lists instructions that ran with a partial warp and local-memory traffic.
"""
import collections
import csv
import re
import sys

STALLS = ["stall_wait", "stall_math", "stall_short_sb", "stall_long_sb", "stall_mio", "stall_no_inst",
          "stall_branch_resolving", "stall_lg", "stall_barrier", "stall_not_selected", "stall_dispatch"]


def kernels(path):
    rows = list(csv.reader(open(path)))
    starts = [i for i, r in enumerate(rows) if r and r[0] == "Kernel Name"]
    out = []
    for n, i in enumerate(starts):
        end = starts[n + 1] if n + 1 < len(starts) else len(rows)
        head = rows[i + 1]
        ix = {name: k for k, name in enumerate(head)}
        body = [r for r in rows[i + 2:end] if len(r) >= len(head)]
        out.append((rows[i][1], ix, body))
    # ncu prints every kernel twice when two views are exported; keep distinct names in order
    seen, uniq = set(), []
    for k in out:
        key = (k[0], len(k[2]))
        if key not in seen:
            seen.add(key)
            uniq.append(k)
    return uniq


def num(r, ix, name):
    v = r[ix[name]] if name in ix else ""
    try:
        return int(float(v)) if v else 0
    except ValueError:
        return 0


def main():
    ks = kernels(sys.argv[1])
    which = int(sys.argv[2]) if len(sys.argv) > 2 else len(ks) - 1
    name, ix, body = ks[which]
    tot = sum(num(r, ix, "# Samples") for r in body)
    inst = sum(num(r, ix, "Instructions Executed") for r in body)
    print(f"{name[:90]}\nsamples {tot}  instructions executed {inst / 1e6:.1f} M  static instructions {len(body)}")
    # dispatch block: the most frequent execution count among the hottest instructions
    counts = collections.Counter(num(r, ix, "Instructions Executed") for r in body)
    per_op = max((c for c, k in counts.items() if k >= 20), default=0)
    disp = [r for r in body if num(r, ix, "Instructions Executed") == per_op]
    ds = sum(num(r, ix, "# Samples") for r in disp)
    print(f"dispatch block: {len(disp)} instructions x {per_op} executions, {ds} samples ({100.0 * ds / max(tot, 1):.1f} %)")
    regions, cur, mix, n = [], collections.Counter(), collections.Counter(), 0
    for k, r in enumerate(body):
        cur["samp"] += num(r, ix, "# Samples")
        cur["ex"] += num(r, ix, "Instructions Executed")
        cur["wf"] += num(r, ix, "L1 Wavefronts Shared")
        cur["exc"] += num(r, ix, "L1 Wavefronts Shared Excessive")
        for s in STALLS:
            cur[s] += num(r, ix, s)
        m = re.match(r"\s*(@!?U?P\d+\s+)?([A-Z0-9_]+)", r[ix["Source"]])
        mix[m.group(2) if m else "?"] += 1
        n += 1
        if "BRX" in r[ix["Source"]] or k == len(body) - 1:
            regions.append((k, n, cur, mix))
            cur, mix, n = collections.Counter(), collections.Counter(), 0
    for k, n, c, mix in regions:
        if c["samp"] == 0:
            continue
        st = sorted(((c[s], s[6:]) for s in STALLS), reverse=True)[:4]
        print(f"  ..{k:6d} n={n:5d} samples {c['samp']:6d} ({100.0 * c['samp'] / max(tot, 1):5.1f} %) "
              f"ex {c['ex'] / 1e6:7.1f} M  smem wf {c['wf'] / 1e6:6.1f} M (+{c['exc'] / 1e6:.1f} M)  "
              + " ".join(f"{s}:{v}" for v, s in st) + "  | " + " ".join(f"{a}:{b}" for a, b in mix.most_common(5)))
    part = [(k, r) for k, r in enumerate(body) if num(r, ix, "Instructions Executed") >= per_op // 8 > 0
            and 0 < float(r[ix["Avg. Threads Executed"]] or 32) < 31.5 and num(r, ix, "# Samples") > 50]
    print(f"instructions run by a partial warp with > 50 samples: {len(part)}")
    for k, r in part[:12]:
        print(f"   {k:6d} {r[ix['Source']][:56]:56s} threads {float(r[ix['Avg. Threads Executed']]):4.1f} "
              f"executed {num(r, ix, 'Instructions Executed')} samples {num(r, ix, '# Samples')}")
    loc = [r for r in body if re.search(r"\b(STL|LDL)", r[ix["Source"]])]
    print(f"local-memory instructions: {len(loc)} static, {sum(num(r, ix, 'Instructions Executed') for r in loc) / 1e6:.1f} M executed, "
          f"{sum(num(r, ix, '# Samples') for r in loc)} samples")


if __name__ == "__main__":
    main()
