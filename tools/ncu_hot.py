#!/usr/bin/env python
"""Top stall lines of one kernel from an ncu report: python tools/ncu_hot.py rep kernel_regex [N]"""
import csv, subprocess, sys, io
rep, kern = sys.argv[1], sys.argv[2]
n = int(sys.argv[3]) if len(sys.argv) > 3 else 25
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--kernel-name", "regex:" + kern, "--launch-skip", "0", "--launch-count", "1"], capture_output=True, text=True).stdout
lines = out.splitlines()
start = next(i for i, l in enumerate(lines) if l.startswith('"Address"'))
end = next((i for i, l in enumerate(lines) if i > start and l.startswith('"Kernel Name"')), len(lines))
rows = list(csv.DictReader(io.StringIO("\n".join(lines[start:end]))))
tot = sum(int(r["# Samples"] or 0) for r in rows)
print("total samples", tot, "instructions", len(rows))
stalls = [k for k in rows[0] if k.startswith("stall_") and "Not Issued" not in k]
agg = {k: sum(int(r[k] or 0) for r in rows) for k in stalls}
print({k: v for k, v in sorted(agg.items(), key=lambda kv: -kv[1]) if v})
for i, r in sorted(enumerate(rows), key=lambda ir: -int(ir[1]["# Samples"] or 0))[:n]:
    top = sorted(((int(r[k] or 0), k) for k in stalls), reverse=True)[:2]
    print(f'{i:5d} {int(r["# Samples"]):6d} {r["Source"].strip()[:70]:70s} {top}')
