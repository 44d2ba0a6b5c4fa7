#!/usr/bin/env python
"""Turn an .ncu-rep into a short markdown summary (key raw metrics, stall-reason shares and the
hottest SASS instructions by sampled stalls).  Usage: ncu_summary.py rep.ncu-rep out.md [title]"""
import csv
import io
import subprocess
import sys


def ncu_csv(rep, page):
    out = subprocess.run(["ncu", "-i", rep, "--page", page, "--csv"], capture_output=True, text=True).stdout
    return list(csv.reader(io.StringIO(out)))


KEYS = ["launch__grid_size", "launch__block_size", "launch__registers_per_thread", "launch__shared_mem_per_block_dynamic",
        "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "lts__t_sector_hit_rate.pct",
        "sm__throughput.avg.pct_of_peak_sustained_elapsed", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "smsp__issue_active.avg.pct_of_peak_sustained_active", "smsp__inst_executed.sum",
        "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active",
        "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum"]


def main():
    rep, out = sys.argv[1], sys.argv[2]
    title = sys.argv[3] if len(sys.argv) > 3 else rep
    raw = ncu_csv(rep, "raw")
    hdr, units = raw[0], raw[1]
    lines = [f"# {title}", "", f"source: `{rep.split('/')[-1]}` (ncu --set full --clock-control none --import-source on, B200)", ""]
    for k, r in enumerate(raw[2:]):
        name = r[hdr.index("Kernel Name")]
        lines += [f"## launch {k}: `{name[:100]}`", "", "| metric | value |", "|---|---|"]
        for key in KEYS:
            if key in hdr:
                i = hdr.index(key)
                lines.append(f"| {key} | {r[i]} {units[i]} |")
        lines.append("")
    src = ncu_csv(rep, "source")
    h = None
    data = []
    for r in src:
        if r and r[0] == "Kernel Name":
            if data:
                break
            continue
        if r and r[0] == "Address":
            h = r
            continue
        if h and len(r) == len(h):
            data.append(r)
    if h and data:
        ci = {n: i for i, n in enumerate(h)}
        stalls = [n for n in h if n.startswith("stall_") and "Not Issued" not in n]
        tot = sum(int(r[ci["# Samples"]] or 0) for r in data) or 1
        agg = sorted(((sum(int(r[ci[s]] or 0) for r in data), s) for s in stalls), reverse=True)
        lines += ["## warp-stall samples (first captured launch)", "", "| stall | share |", "|---|---|"]
        lines += [f"| {s} | {100 * v / tot:.1f}% |" for v, s in agg[:8]]
        lines += ["", "## hottest SASS instructions by samples", "", "| samples | SASS | dominant stall | executed |", "|---|---|---|---|"]
        for r in sorted(data, key=lambda r: -int(r[ci["# Samples"]] or 0))[:14]:
            ss = {s: int(r[ci[s]] or 0) for s in stalls}
            top = max(ss, key=ss.get)
            lines.append(f"| {r[ci['# Samples']]} | `{r[ci['Source']][:70]}` | {top} | {r[ci['Instructions Executed']]} |")
    open(out, "w").write("\n".join(lines) + "\n")
    print("wrote", out)


if __name__ == "__main__":
    main()
