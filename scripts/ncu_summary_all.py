#!/usr/bin/env python
"""One table over every kernel of an `ncu --set full` report: for the LAST captured launch of each distinct kernel
(the driver launches each twice; the second is warm) — duration, DRAM bytes read / written, achieved DRAM GB/s and its
fraction of the measured copy peak (MEASURED_PEAKS.json), registers, issue-active %, and the two largest warp-stall
reasons from the source page.  Usage: ncu_summary_all.py rep.ncu-rep out.md [title]"""
import collections
import csv
import io
import json
import re
import subprocess
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent


def ncu_csv(rep, page):
    out = subprocess.run(["ncu", "-i", rep, "--page", page, "--csv"], capture_output=True, text=True).stdout
    return list(csv.reader(io.StringIO(out)))


def num(x):
    try:
        return float(str(x).replace(",", ""))
    except ValueError:
        return 0.0


def to_bytes(v, unit):
    return num(v) * {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}.get(unit, 1)


def to_us(v, unit):
    return num(v) * {"ns": 1e-3, "us": 1, "ms": 1e3, "s": 1e6}.get(unit, 1)


def short(name):
    name = re.sub(r"^void ", "", name)
    name = re.sub(r"\(.*$", "", name)
    return name.replace("dlion::", "").replace("__nv_bfloat16", "bf16")[:96]


def stall_shares(rep):
    """kernel id order -> {stall: share} from the source page (one block per captured launch)."""
    rows = ncu_csv(rep, "source")
    out, h, cur = [], None, None
    for r in rows:
        if r and r[0] == "Kernel Name":
            continue
        if r and r[0] == "Address":
            h = r
            cur = collections.Counter()
            out.append(cur)
            continue
        if h and cur is not None and len(r) == len(h):
            for i, n in enumerate(h):
                if n.startswith("stall_") and "Not Issued" not in n:
                    cur[n] += int(num(r[i]))
    return out


def main():
    rep, out = sys.argv[1], sys.argv[2]
    title = sys.argv[3] if len(sys.argv) > 3 else rep
    peak = 6545.0
    f = ROOT / "MEASURED_PEAKS.json"
    if f.exists():
        peak = float(json.loads(f.read_text())["hbm_gbs"])
    raw = ncu_csv(rep, "raw")
    hdr, units = raw[0], raw[1]
    col = {n: i for i, n in enumerate(hdr)}
    stalls = stall_shares(rep)
    last = collections.OrderedDict()
    for k, r in enumerate(raw[2:]):
        last[short(r[col["Kernel Name"]])] = (k, r)

    def get(r, key):
        i = col.get(key)
        return (r[i], units[i]) if i is not None else ("0", "")

    lines = [f"# {title}", "", f"source: `{Path(rep).name}` — `ncu --set full --clock-control none --import-source on` on one B200, "
             "driver `bench/profile_all_kernels.py` (every kernel at its training-path shape, second launch of each).",
             f"Roofline denominator: measured copy peak {peak:.0f} GB/s (MEASURED_PEAKS.json).  Durations under ncu are "
             "serialised and cold-L2; compare fractions, not absolutes.", "",
             "| kernel | us | DRAM read MB | DRAM write MB | GB/s | of copy peak | regs | issue active % | top stalls |",
             "|---|---:|---:|---:|---:|---:|---:|---:|---|"]
    for name, (k, r) in last.items():
        us = to_us(*get(r, "gpu__time_duration.sum"))
        rd = to_bytes(*get(r, "dram__bytes_read.sum"))
        wr = to_bytes(*get(r, "dram__bytes_write.sum"))
        gbs = (rd + wr) / us / 1e3 if us else 0.0
        regs = get(r, "launch__registers_per_thread")[0]
        issue = num(get(r, "smsp__issue_active.avg.pct_of_peak_sustained_active")[0])
        st = stalls[k] if k < len(stalls) else {}
        tot = sum(st.values()) or 1
        top = ", ".join(f"{s.replace('stall_', '')} {100 * v / tot:.0f}%" for s, v in collections.Counter(st).most_common(2))
        lines.append(f"| `{name}` | {us:.1f} | {rd / 1e6:.1f} | {wr / 1e6:.1f} | {gbs:.0f} | {gbs / peak:.2f} | {regs} | {issue:.0f} | {top} |")
    Path(out).write_text("\n".join(lines) + "\n")
    print("wrote", out, len(last), "kernels")


if __name__ == "__main__":
    main()
