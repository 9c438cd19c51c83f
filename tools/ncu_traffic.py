"""DRAM traffic per frame of each kernel from an ncu --set full report of ONE extraction step:
   python tools/ncu_traffic.py report.ncu-rep frames_in_the_profiled_step out.json
Launches of the same kernel inside the step (the 7 pyramid levels) are summed; if a kernel shows up more often
than once per step (the capture window wrapped around), its launches are averaged per occurrence of the step."""
import collections
import csv
import json
import re
import subprocess
import sys

rep, frames, out = sys.argv[1], int(sys.argv[2]), sys.argv[3]
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(raw.splitlines()))
H, U = rows[0], rows[1]
ik, ir, iw = H.index("Kernel Name"), H.index("dram__bytes_read.sum"), H.index("dram__bytes_write.sum")
it = H.index("gpu__time_duration.sum")
scale = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
per_step = {"k_resize": 7}
agg = collections.OrderedDict()
for r in rows[2:]:
    if len(r) < len(H):
        continue
    k = re.sub(r"^void |ygz::|<.*$", "", r[ik].split("(")[0])   # template instantiations of one kernel count together
    a = agg.setdefault(k, [0, 0.0, 0.0, 0.0])
    a[0] += 1
    a[1] += float(r[ir]) * scale[U[ir]]
    a[2] += float(r[iw]) * scale[U[iw]]
    a[3] += float(r[it]) * {"ns": 1e-3, "us": 1.0, "ms": 1e3}.get(U[it], 1.0)
res = {"report": rep.split("/")[-1], "frames": frames, "kernels": {}}
for k, (n, rd, wr, us) in agg.items():
    steps = n / per_step.get(k, 1)
    res["kernels"][k] = {"launches_per_step": per_step.get(k, 1), "dram_read_bytes_per_frame": rd / steps / frames,
                         "dram_write_bytes_per_frame": wr / steps / frames, "ncu_us_per_step": us / steps}
json.dump(res, open(out, "w"), indent=1)
print(json.dumps(res, indent=1))
