#!/usr/bin/env python3
"""Where the instructions of a kernel go, from an `ncu --set full --import-source on` report (needs -lineinfo):

  python tools/ncu_hot_code.py gpurun_out/rollout.ncu-rep profiles/r2p_rollout_hot_code.txt ["command that was profiled"]

Writes (1) the hot-code footprint: how many 128-byte instruction-cache lines hold 50 / 90 / 99 % of the executed warp
instructions (the playout kernel is bound by instruction fetch, so this is its working set against the 32 KB L1.5 / 6 KB L0
instruction caches), (2) executed warp instructions and active lanes by source file, (3) the source lines that execute the
most warp instructions, (4) every run-time division sequence (MUFU.RCP) with its share. Run here: the CPU box has ncu."""
import csv
import io
import os
import subprocess
import sys

rep, out = sys.argv[1], sys.argv[2]
cmd = sys.argv[3] if len(sys.argv) > 3 else ""


def page(extra):
    txt = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"] + extra, capture_output=True, text=True, check=True).stdout
    return list(csv.reader(io.StringIO(txt)))


lines = []
w = lines.append

# ---- SASS view: footprint
rows = page([])
kernel = rows[0][1]
hdr, data = rows[1], rows[2:]
ix = {h: i for i, h in enumerate(hdr)}
addr = [int(r[ix["Address"]], 16) for r in data]
ex = [int(r[ix["Instructions Executed"]]) for r in data]
th = [int(r[ix["Thread Instructions Executed"]]) for r in data]
smp = [int(r[ix["# Samples"]]) for r in data]
noi = [int(r[ix["stall_no_inst"]]) for r in data]
tot = sum(ex)
w(f"kernel: {kernel}")
if cmd:
    w(f"command: {cmd}")
w(f"SASS instructions {len(data)} = {len(data) * 16 / 1024:.1f} KB; executed at least once {sum(1 for e in ex if e)} = "
  f"{sum(1 for e in ex if e) * 16 / 1024:.1f} KB")
w(f"executed warp instructions {tot}; active lanes per instruction {sum(th) / tot:.2f}; "
  f"stall samples {sum(smp)}, no_instruction {sum(noi) / max(sum(smp), 1) * 100:.1f} %")
per_line = {}
for a, e in zip(addr, ex):
    per_line[(a - addr[0]) // 128] = per_line.get((a - addr[0]) // 128, 0) + e
acc, marks, mi = 0, [0.5, 0.7, 0.8, 0.9, 0.95, 0.99], 0
for k, e in enumerate(sorted(per_line.values(), reverse=True)):
    acc += e
    while mi < len(marks) and acc >= marks[mi] * tot:
        w(f"  {int(marks[mi] * 100):2d} % of the executed warp instructions lie in {k + 1:3d} lines of 128 B = {(k + 1) / 8:.1f} KB")
        mi += 1

# ---- CUDA + SASS view: by file, by line, divisions
rows = page(["--print-source", "cuda,sass"])
cur, hdr, src, line, divs = None, None, [], None, []
for r in rows:
    if r and r[0] == "File Path":
        cur = os.path.basename(r[1])
    elif r and r[0] == "Line No":
        hdr = r
        iE, iT = hdr.index("Instructions Executed"), hdr.index("Thread Instructions Executed")
    elif hdr and len(r) == len(hdr):
        if r[0] != "":
            line = (cur, r[0], r[1].strip())
            src.append((cur, r))
        elif "MUFU.RCP" in r[3]:
            divs.append((int(r[iE]), line))
stot = sum(int(r[iE]) for _, r in src)
w("")
w("by source file: share of executed warp instructions, active lanes")
byf = {}
for f, r in src:
    d = byf.setdefault(f, [0, 0])
    d[0] += int(r[iE])
    d[1] += int(r[iT])
for f, d in sorted(byf.items(), key=lambda kv: -kv[1][0]):
    if d[0] * 1000 >= stot:
        w(f"  {f:32s} {d[0] / stot * 100:5.1f} %  lanes {d[1] / max(d[0], 1):5.1f}")
w("")
w("source lines by executed warp instructions (share, cumulative, active lanes)")
src.sort(key=lambda fr: -int(fr[1][iE]))
acc = 0
for f, r in src[:40]:
    e, t = int(r[iE]), int(r[iT])
    acc += e
    w(f"  {e / stot * 100:5.2f} % ({acc / stot * 100:4.1f}) lanes {t / max(e, 1):5.1f}  {f}:{r[0]}  {r[1].strip()[:100]}")
w("")
w("run-time division sequences (one MUFU.RCP each; a sequence is ~20 instructions, a 64-bit one ~55): executions as a share of all")
for e, l in sorted(divs, key=lambda x: -x[0])[:12]:
    w(f"  {e / tot * 100:6.3f} %  {l[0]}:{l[1]}  {l[2][:100]}")
open(out, "w").write("\n".join(lines) + "\n")
print("\n".join(lines))
