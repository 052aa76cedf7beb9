"""Per-source-line instruction and stall-sample totals of one kernel from an ncu report (development helper).

  python tools/ncu_lines.py REPORT.ncu-rep KERNEL_REGEX[::MANGLED_REGEX] [launch_skip] [top_n]

ncu's CSV source page is SASS-only, so the SASS rows (in program order) are joined with the line table that
`nvdisasm -g` prints for the cubin embedded in the in-tree library (compiled with -lineinfo).
"""
import csv
import io
import os
import re
import subprocess
import sys
import tempfile
from collections import defaultdict

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.environ.get("NCU_LIB") or os.path.join(ROOT, "pathtrace_b200", "lib", "libpathtrace_b200.so")


def line_table(kernel_re):
    """[(offset, 'file:line')] for the first function whose mangled name matches kernel_re."""
    tmp = tempfile.mkdtemp(prefix="ncu_lines_")
    subprocess.check_call(["cuobjdump", "-xelf", "all", LIB], cwd=tmp, stdout=subprocess.DEVNULL)
    for cub in sorted(os.listdir(tmp)):
        dis = subprocess.run(["nvdisasm", "-g", "-c", os.path.join(tmp, cub)], capture_output=True, text=True).stdout
        cur, where, out = None, "?", []
        for ln in dis.splitlines():
            m = re.match(r"\.text\.(\S+):", ln)
            if m:
                if out:
                    return out
                cur = m.group(1) if re.search(kernel_re, m.group(1)) else None
                continue
            if cur is None:
                continue
            m = re.search(r'//## File "([^"]+)", line (\d+)', ln)
            if m:
                where = f"{os.path.basename(m.group(1))}:{m.group(2)}"
                continue
            m = re.match(r"\s+/\*([0-9a-f]{4,})\*/\s+(.*?);", ln)
            if m:
                out.append((int(m.group(1), 16), where, m.group(2).strip()))
        if out:
            return out
    raise SystemExit(f"no function matching {kernel_re}")


def main():
    rep, kre = sys.argv[1], sys.argv[2]
    nre = kre                      # "ncu_regex::mangled_regex" when the demangled and mangled names need different patterns
    if "::" in kre:
        nre, kre = kre.split("::", 1)
    skip = sys.argv[3] if len(sys.argv) > 3 else "0"
    top = int(sys.argv[4]) if len(sys.argv) > 4 else 40
    txt = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--kernel-name", f"regex:{nre}", "--launch-skip", skip,
                          "--launch-count", "1"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(txt)))
    hdr_i = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
    hdr = rows[hdr_i]
    body = []
    for r in rows[hdr_i + 1:]:          # first matching launch only
        if len(r) != len(hdr) or r[0] == "Address":
            break
        body.append(r)
    ci, cs, ct = hdr.index("Instructions Executed"), hdr.index("# Samples"), hdr.index("Thread Instructions Executed")
    stall_cols = [i for i, h in enumerate(hdr) if h.startswith("stall_") and "Not Issued" not in h]
    tab = line_table(kre)
    if len(tab) != len(body):
        print(f"warning: {len(tab)} SASS lines in the cubin vs {len(body)} in the report (stale library?)", file=sys.stderr)
    base = int(body[0][0], 16)
    off2line = {o: w for o, w, _ in tab}
    inst, samp, thr = defaultdict(int), defaultdict(int), defaultdict(int)
    stalls = defaultdict(lambda: defaultdict(int))
    for r in body:
        w = off2line.get(int(r[0], 16) - base, "?")
        inst[w] += int(r[ci]); samp[w] += int(r[cs]); thr[w] += int(r[ct])
        for i in stall_cols:
            if r[i] not in ("", "0"):
                stalls[w][hdr[i]] += int(r[i])
    if os.environ.get("OPS"):        # OPS=1: executed-instruction histogram by opcode (and by opcode within the top lines)
        hs = hdr.index("Source")
        ops, ops_line = defaultdict(int), defaultdict(lambda: defaultdict(int))
        for r in body:
            toks = [t for t in r[hs].split() if not t.startswith("@")]
            op = toks[0].rstrip(";") if toks else "?"
            op = ".".join(op.split(".")[:2]) if op.startswith(("LDS", "STS", "MUFU", "ATOMS", "RED", "LDG", "STG")) else op.split(".")[0]
            ops[op] += int(r[ci])
            ops_line[off2line.get(int(r[0], 16) - base, "?")][op] += int(r[ci])
        tot = sum(ops.values())
        print(f"{tot} warp instructions")
        for op, v in sorted(ops.items(), key=lambda kv: -kv[1])[:top]:
            print(f"  {op:14s} {100 * v / tot:6.2f} %")
        for w in sorted(inst, key=lambda k: -inst[k])[:int(os.environ.get("OPS_LINES", "12"))]:
            print(f"{w:28s} {100 * inst[w] / tot:6.2f} %  " + " ".join(f"{o}={100 * v / tot:.2f}" for o, v in sorted(ops_line[w].items(), key=lambda kv: -kv[1])[:8]))
        return
    ti, ts = sum(inst.values()), sum(samp.values())
    print(f"kernel {kre}: {ti} warp instructions, {ts} samples, avg threads {sum(thr.values()) / max(ti, 1):.2f}")
    print(f"{'line':28s} {'inst%':>6s} {'samp%':>6s} {'thr':>5s}  top stalls")
    for w in sorted(inst, key=lambda k: -samp[k])[:top]:
        st = sorted(stalls[w].items(), key=lambda kv: -kv[1])[:3]
        print(f"{w:28s} {100 * inst[w] / ti:6.2f} {100 * samp[w] / max(ts, 1):6.2f} {thr[w] / max(inst[w], 1):5.1f}  " +
              " ".join(f"{k[6:]}={v}" for k, v in st))


if __name__ == "__main__":
    main()
