"""Summarises an .ncu-rep (ncu --set full) and/or a launch list csv into markdown + the
roofline traffic json bench.py reads.
    python tools/ncu_summary.py --rep gpurun_out/prof.ncu-rep --launches gpurun_out/launches.csv \
        --out profiles/r1_bench_ncu.md --traffic profiles/roofline_traffic.json"""
import argparse
import collections
import csv
import io
import json
import subprocess

METRICS = [
    "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
    "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "lts__throughput.avg.pct_of_peak_sustained_elapsed",
    "l1tex__throughput.avg.pct_of_peak_sustained_elapsed", "smsp__inst_executed.sum",
    "smsp__issue_active.avg.pct_of_peak_sustained_active", "sm__warps_active.avg.pct_of_peak_sustained_active",
    "smsp__thread_inst_executed_per_inst_executed.ratio", "launch__registers_per_thread", "launch__grid_size",
    "launch__block_size", "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active",
]
UNIT_BYTES = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}


def short_name(full):
    """`void dirt::<unnamed>::k_setup<1>(dirt::FwdParams, int)` -> `k_setup<1>`; library kernels keep their name."""
    base = full.split("(")[0].replace("void ", "")
    i = base.find("k_")
    return base[i:] if i >= 0 and not base.startswith("at::") else base


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--rep")
    ap.add_argument("--launches")
    ap.add_argument("--out", required=True)
    ap.add_argument("--traffic")
    ap.add_argument("--title", default="ncu summary")
    a = ap.parse_args()
    lines = ["# " + a.title, ""]
    traffic = {}
    if a.launches:
        tot = collections.OrderedDict()
        for r in csv.DictReader(l for l in open(a.launches) if l.startswith('"')):
            if r.get("Metric Name") != "gpu__time_duration.sum":
                continue
            k = short_name(r["Kernel Name"])
            t = tot.setdefault(k, [0, 0.0])
            t[0] += 1
            t[1] += float(r["Metric Value"].replace(",", "")) * {"ns": 1e-3, "us": 1.0, "ms": 1e3}.get(r["Metric Unit"], 1e-3)
        total = sum(v[1] for v in tot.values()) or 1.0
        lines += ["## Launch list (`ncu --metrics gpu__time_duration.sum --clock-control none`; cold-cache, serialised: shares matter)", "",
                  "| kernel | launches | total us | mean us | share |", "|---|---:|---:|---:|---:|"]
        for k, (n, us) in sorted(tot.items(), key=lambda kv: -kv[1][1]):
            lines.append("| `%s` | %d | %.1f | %.1f | %.1f%% |" % (k, n, us, us / n, 100 * us / total))
        lines.append("")
    if a.rep:
        raw = subprocess.run(["ncu", "-i", a.rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
        rows = list(csv.reader(io.StringIO(raw)))
        hdr, units = rows[0], rows[1]
        lines += ["## `ncu --set full --clock-control none` (the launches of one step)", ""]
        for r in rows[2:]:
            name = short_name(r[hdr.index("Kernel Name")])
            lines += ["### `%s`" % name, "", "| metric | value | unit |", "|---|---:|---|"]
            rd = wr = 0.0
            for m in METRICS:
                if m in hdr:
                    i = hdr.index(m)
                    lines.append("| %s | %s | %s |" % (m, r[i], units[i]))
                    if m == "dram__bytes_read.sum":
                        rd = float(r[i].replace(",", "")) * UNIT_BYTES.get(units[i], 1)
                    if m == "dram__bytes_write.sum":
                        wr = float(r[i].replace(",", "")) * UNIT_BYTES.get(units[i], 1)
            key = "k_" + name.split("<")[0].replace("k_", "")
            traffic[key] = traffic.get(key, 0.0) + rd + wr          # capture one step: launches of a kernel add up
            # what actually binds the kernel: issue slots, the LSU data pipe, DRAM (per launch, in capture order)
            pipes = traffic.setdefault("_pipes", {}).setdefault(key, [])
            entry = {}
            for label, m in (("issue_active_pct", "smsp__issue_active.avg.pct_of_peak_sustained_active"),
                             ("lsu_data_pipe_pct", "l1tex__data_pipe_lsu_wavefronts.avg.pct_of_peak_sustained_elapsed"),
                             ("dram_pct", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed"),
                             ("duration_us", "gpu__time_duration.sum")):
                if m in hdr:
                    try:
                        entry[label] = round(float(r[hdr.index(m)].replace(",", "")), 1)
                    except ValueError:
                        pass
            pipes.append(entry)
            lines.append("")
    open(a.out, "w").write("\n".join(lines) + "\n")
    if a.traffic:
        json.dump(traffic, open(a.traffic, "w"), indent=1)
    print("\n".join(lines))


if __name__ == "__main__":
    main()
