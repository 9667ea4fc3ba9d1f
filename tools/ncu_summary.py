#!/usr/bin/env python
"""Summarise ncu outputs into a small tracked text file.
  python tools/ncu_summary.py --launches gpurun_out/launches.csv --rep gpurun_out/prof.ncu-rep --out profiles/NAME.md --title "..."
"""
import argparse, collections, csv, io, statistics, subprocess

KEYS = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
        "launch__registers_per_thread", "launch__grid_size", "launch__block_size", "launch__waves_per_multiprocessor",
        "launch__occupancy_limit_registers", "launch__occupancy_limit_shared_mem", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "sm__throughput.avg.pct_of_peak_sustained_elapsed", "smsp__cycles_active.avg", "lts__t_sectors_srcunit_tex_op_read.sum",
        "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active"]


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--launches"), ap.add_argument("--rep"), ap.add_argument("--out", required=True), ap.add_argument("--title", default="ncu summary")
    ap.add_argument("--note", default="")
    a = ap.parse_args()
    out = [f"# {a.title}", "", a.note, ""]
    if a.launches:
        rows = list(csv.DictReader(l for l in open(a.launches) if l.startswith('"')))
        d = collections.OrderedDict()
        for r in rows:
            d.setdefault((r["Kernel Name"].split("(")[0], r["Grid Size"], r["Block Size"]), []).append(float(r["Metric Value"]))
        tot = sum(sum(v) for v in d.values())
        out += ["## Launch list (`ncu --metrics gpu__time_duration.sum --clock-control none`; cold-cache, serialised: compare SHARES)", "",
                "| kernel | grid | block | launches | median ns | min | max | share of listed time |", "|---|---|---|---|---|---|---|---|"]
        for (k, g, b), v in d.items():
            out.append(f"| `{k}` | {g} | {b} | {len(v)} | {statistics.median(v):.0f} | {min(v):.0f} | {max(v):.0f} | {sum(v) / tot:.1%} |")
        out.append("")
    if a.rep:
        txt = subprocess.run(["ncu", "-i", a.rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
        rows = list(csv.reader(io.StringIO(txt)))
        hdr, units = rows[0], rows[1]
        out += ["## `ncu --set full` capture (per launch)", ""]
        seen = set()
        for r in rows[2:]:
            name = r[hdr.index("Kernel Name")].split("(")[0]
            if name in seen:
                continue
            seen.add(name)
            out.append(f"### `{name}`")
            for k in KEYS:
                if k in hdr:
                    i = hdr.index(k)
                    out.append(f"- {k} = {r[i]} {units[i]}")
            if "dram__bytes_read.sum" in hdr:
                def val(k):
                    i = hdr.index(k)
                    x = float(r[i].replace(",", ""))
                    u = units[i]
                    return x * {"Mbyte": 1e6, "Kbyte": 1e3, "Gbyte": 1e9, "byte": 1}.get(u, 1)
                out.append(f"- **traffic** (dram read+write) = {val('dram__bytes_read.sum') + val('dram__bytes_write.sum'):.0f} B")
            out.append("")
    open(a.out, "w").write("\n".join(out) + "\n")
    print("\n".join(out))


if __name__ == "__main__":
    main()
