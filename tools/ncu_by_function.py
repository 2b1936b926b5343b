#!/usr/bin/env python
"""Per-function share of warp-state samples and executed instructions of one kernel in an .ncu-rep
(needs -lineinfo and --import-source on):   python tools/ncu_by_function.py report.ncu-rep [out.txt]
Inlined code is attributed to the source line it came from; a line belongs to the last function definition above it."""
import csv, io, re, subprocess, sys, collections

rep = sys.argv[1]
txt = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass,cuda"], capture_output=True, text=True).stdout
blocks, cur = [], None
for row in csv.reader(io.StringIO(txt)):
    if len(row) == 2 and row[0] == "File Path":
        cur = dict(path=row[1], rows=[], head=None)
        blocks.append(cur)
    elif cur is not None and row and row[0] == "Line No":
        cur["head"] = row
    elif cur is not None and cur["head"] and len(row) == len(cur["head"]):
        cur["rows"].append(row)
DEF = re.compile(r"^\s*(?:template\s*<[^>]*>\s*)?(?:SRT_HD\w*|__device__|__global__|static|inline)\b.*?\b((?:srt|team)_\w+)\s*\(")
per_fn = collections.defaultdict(lambda: [0, 0, collections.Counter()])
tot_s = tot_i = 0
for b in blocks:
    h = b["head"]
    isrc, isamp, iex = 1, h.index("# Samples"), h.index("Instructions Executed")
    stalls = [(k, c) for k, c in enumerate(h) if c.startswith("stall_") and "Not Issued" not in c]
    # function of every line of the file, from the file itself (definition lines carry no instructions of their own)
    fn_of, fn = {}, "(top of %s)" % b["path"].split("/")[-1]
    try:
        for ln, text in enumerate(open(b["path"], errors="replace"), 1):
            m = DEF.match(text)
            if m and not text.rstrip().endswith(";"):
                fn = m.group(1)
            fn_of[ln] = fn
    except OSError:
        pass
    src_rows = [r for r in b["rows"] if r[0].strip().isdigit()]  # rows without a line number are the SASS of the line above
    for r in sorted(src_rows, key=lambda r: int(r[0])):
        fn = fn_of.get(int(r[0]), fn)
        s, e = int(r[isamp] or 0), int(r[iex] or 0)
        key = "%s:%s" % (b["path"].split("/")[-1], fn)
        per_fn[key][0] += s
        per_fn[key][1] += e
        for k, c in stalls:
            per_fn[key][2][c] += int(r[k] or 0)
        tot_s += s
        tot_i += e
lines = ["%-52s %8s %7s %12s %7s  top stall reasons" % ("file:function", "samples", "%", "warp-instr", "%")]
for key, (s, e, st) in sorted(per_fn.items(), key=lambda kv: -kv[1][0]):
    if s < 0.003 * tot_s and e < 0.003 * tot_i:
        continue
    top = ", ".join("%s %.0f%%" % (k.replace("stall_", ""), 100.0 * v / max(s, 1)) for k, v in st.most_common(3))
    lines.append("%-52s %8d %6.1f%% %12d %6.1f%%  %s" % (key, s, 100.0 * s / max(tot_s, 1), e, 100.0 * e / max(tot_i, 1), top))
lines.append("total samples %d, warp-level instructions executed %d" % (tot_s, tot_i))
out = "\n".join(lines)
print(out)
if len(sys.argv) > 2:
    open(sys.argv[2], "w").write(out + "\n")
