"""Executed-instruction shares of a kernel split at its barriers (phases) and, inside each phase, by source function
(development helper; same inputs as tools/ncu_lines.py).

  python tools/ncu_phases.py REPORT.ncu-rep NCU_REGEX::MANGLED_REGEX
"""
import csv
import io
import subprocess
import sys
from collections import defaultdict

import ncu_lines

# (file, first line, last line, label): pt_wave.cuh / pt_math.cuh functions of interest
def _ranges():
    import os
    import re
    out = []
    csrc = os.path.join(ncu_lines.ROOT, "pathtrace_b200", "csrc")
    for fn in ("pt_wave.cuh", "pt_math.cuh", "pt_shade.cuh"):
        starts = []
        for i, ln in enumerate(open(os.path.join(csrc, fn)), 1):
            m = re.match(r"(?:PT_DEV|template|inline|__global__)[^(]*?\b(\w+)\s*\(", ln)
            if m and not ln.startswith("template <"):
                starts.append((i, m.group(1)))
        for (a, name), (b, _) in zip(starts, starts[1:] + [(10 ** 9, "")]):
            out.append((fn, a, b - 1, name))
    return out


def main():
    rep, kre = sys.argv[1], sys.argv[2]
    nre, kre = kre.split("::", 1) if "::" in kre else (kre, kre)
    txt = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--kernel-name", f"regex:{nre}", "--launch-count", "1"],
                         capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(txt)))
    hi = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
    hdr = rows[hi]
    body = []
    for r in rows[hi + 1:]:             # first matching launch only
        if len(r) != len(hdr) or r[0] == "Address":
            break
        body.append(r)
    ci, hs, ct, cs = hdr.index("Instructions Executed"), hdr.index("Source"), hdr.index("Thread Instructions Executed"), hdr.index("# Samples")
    tab = ncu_lines.line_table(kre)
    if len(tab) != len(body):
        print(f"warning: {len(tab)} SASS lines in the cubin vs {len(body)} in the report (stale library?)", file=sys.stderr)
    rng = _ranges()

    def label(w):
        f, _, l = w.partition(":")
        l = int(l) if l.isdigit() else 0
        for fn, a, b, name in rng:
            if f == fn and a <= l <= b:
                return f"{fn.split('.')[0]}:{name}"
        return f
    tot = sum(int(r[ci]) for r in body)
    tots = sum(int(r[cs]) for r in body) or 1
    seg, segs = defaultdict(lambda: [0, 0, 0]), []
    acc = [0, 0, 0]
    for k, r in enumerate(body):
        w = tab[k][1] if k < len(tab) else "?"
        v = (int(r[ci]), int(r[cs]), int(r[ct]))
        x = seg[label(w)]
        for i in range(3):
            x[i] += v[i]; acc[i] += v[i]
        if "BAR" in r[hs] or k == len(body) - 1:
            segs.append((k, acc, seg))
            seg, acc = defaultdict(lambda: [0, 0, 0]), [0, 0, 0]
    print(f"{tot} warp instructions, {tots} samples")
    for k, a, sg in segs:
        print(f"-- up to SASS line {k}: inst {100 * a[0] / tot:6.2f} %  samples {100 * a[1] / tots:6.2f} %  threads/inst {a[2] / max(a[0], 1):5.1f}")
        if a[0] / tot < 0.02:
            continue
        for name, v in sorted(sg.items(), key=lambda kv: -kv[1][0])[:16]:
            print(f"     {name:34s} inst {100 * v[0] / tot:6.2f} %  samp {100 * v[1] / tots:6.2f} %  thr {v[2] / max(v[0], 1):5.1f}")


if __name__ == "__main__":
    main()
