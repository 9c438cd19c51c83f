"""Per-segment instruction share of one kernel in an .ncu-rep (needs --import-source on / --set full).

    python tools/ncu_lines.py report.ncu-rep kernel_regex [--detail]

The SASS listing is cut into segments at every barrier and at every loop (a backward branch closes a
segment that starts at its target).  For each segment: warp-instructions executed, share of the
kernel, stall samples, and the instructions-per-iteration of loops (executed / executions of the
backward branch).  --detail also prints the listing with per-instruction counts.
"""
import csv
import re
import subprocess
import sys

rep, kern = sys.argv[1], sys.argv[2]
detail = "--detail" in sys.argv
raw = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--kernel-name", "regex:" + kern],
                     capture_output=True, text=True).stdout
rows = list(csv.reader(raw.splitlines()))
first = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
H = rows[first]
iA, iS, iN, iX = H.index("Address"), H.index("Source"), H.index("# Samples"), H.index("Instructions Executed")
ins = []
for r in rows[first + 1:]:
    if not r or r[0] == "Address" or r[0] == "Kernel Name":
        break   # first matching launch only
    ins.append((int(r[iA], 16), r[iS].strip(), int(r[iN] or 0), int(r[iX] or 0)))
base = ins[0][0]
ins = [(a - base, s, n, x) for a, s, n, x in ins]
addr2idx = {a: i for i, (a, _, _, _) in enumerate(ins)}
total = sum(x for _, _, _, x in ins) or 1
samples = sum(n for _, _, n, _ in ins) or 1

# loops: backward branches
loops = []
for i, (a, s, n, x) in enumerate(ins):
    m = re.search(r"BRA(?:\.U)?(?:\.DIV)?\s+(?:U?R?\d+,\s*)?0x([0-9a-f]+)", s)
    if m:
        t = int(m.group(1), 16)
        t = t - base if t >= base else t
        if t <= a and t in addr2idx:
            loops.append((addr2idx[t], i))
cuts = {0, len(ins)}
for b, e in loops:
    cuts.add(b); cuts.add(e + 1)
for i, (a, s, n, x) in enumerate(ins):
    if "BAR." in s:
        cuts.add(i + 1)
cuts = sorted(cuts)
print(f"{kern}: {total/1e6:.1f} M warp-instructions, {len(ins)} SASS lines, {samples} samples")
print(f"{'segment':>15} {'lines':>6} {'Minst':>9} {'inst%':>6} {'smpl%':>6}  note")
for b, e in zip(cuts, cuts[1:]):
    seg = ins[b:e]
    x = sum(t[3] for t in seg); n = sum(t[2] for t in seg)
    note = ""
    for lb, le in loops:
        if lb == b and le == e - 1:
            trips = ins[le][3]
            note = f"loop, {x / max(trips, 1):.1f} inst/iteration, {trips} warp-iterations"
    if "BAR." in seg[-1][1]:
        note += " |barrier"
    print(f"{seg[0][0]:#7x}-{seg[-1][0]:#7x} {len(seg):6d} {x/1e6:9.2f} {100*x/total:6.1f} {100*n/samples:6.1f}  {note}")
if detail:
    for a, s, n, x in ins:
        print(f"{a:#7x} {x:10d} {n:6d}  {s}")
