#!/usr/bin/env python
"""profiles/traffic.json entry from an ncu capture of tools/profile_case.py / tools/c3_age.py:
    python tools/traffic_from_ncu.py WORKLOAD report.ncu-rep plain.log
plain.log is the output of the same command without ncu (accepted steps and kernel geometry of the launch)."""
import csv, io, json, os, re, subprocess, sys
name, rep, plain = sys.argv[1:4]
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
h, d = rows[0], rows[2]
def metric(n):
    i = h.index(n); u = rows[1][i]; v = float(d[i].replace(",", ""))
    return v * {"Gbyte": 1e9, "Mbyte": 1e6, "Kbyte": 1e3, "byte": 1.0, "Tbyte": 1e12}[u]
dram = metric("dram__bytes_read.sum") + metric("dram__bytes_write.sum")
txt = open(plain).read()
acc = int(re.search(r"accepted (\d+)", txt).group(1))
ki = eval(re.search(r"(\{'mode'.*?\})", txt).group(1))
path = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "profiles", "traffic.json")
t = json.load(open(path)) if os.path.exists(path) else {}
t[name] = {"kernel": ki, "dram_bytes_per_accepted_step": dram / acc, "dram_bytes_capture": dram, "accepted_steps_capture": acc,
           "source": os.path.basename(rep), "kernel_name": d[h.index("Kernel Name")] if "Kernel Name" in h else ""}
json.dump(t, open(path, "w"), indent=1)
print(name, t[name])
