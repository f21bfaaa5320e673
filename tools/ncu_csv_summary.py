"""Summaries of the two ncu outputs gpu_round_end.sh brings back (CSV, no .ncu-rep needed):
  python tools/ncu_csv_summary.py raw <raw.csv>          one line per launch: time, dram bytes, instructions, issue %, warps %, registers
  python tools/ncu_csv_summary.py launches <launches.csv>  share of the device time per kernel (gpu__time_duration.sum launch list)"""
import csv
import re
import sys
from collections import OrderedDict


def num(v):
    try:
        return float(v.replace(",", ""))
    except ValueError:
        return float("nan")


def short(name):
    m = re.search(r"(k_[a-z0-9_]+)(<[^>(]*>?)?", name)
    return (m.group(1) + (m.group(2) or "")) if m else name[:60]


def to_bytes(v, unit):
    return num(v) * {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}.get(unit, 1)


def to_us(v, unit):
    return num(v) * {"ns": 1e-3, "us": 1, "usecond": 1, "ms": 1e3, "msecond": 1e3, "nsecond": 1e-3, "second": 1e6, "s": 1e6}.get(unit, 1)


def raw(path):
    rows = [r for r in csv.reader(open(path)) if r]
    i0 = next(i for i, r in enumerate(rows) if "Kernel Name" in r)
    h, u = rows[i0], rows[i0 + 1]
    col = {n: i for i, n in enumerate(h)}

    def get(r, n, conv=None):
        if n not in col:
            return float("nan")
        return conv(r[col[n]], u[col[n]]) if conv else num(r[col[n]])
    print("%-44s %9s %9s %9s %10s %7s %7s %5s %9s" % ("kernel", "time us", "dram rd MB", "dram wr MB", "warp inst", "issue%", "warps%", "regs", "grid"))
    for r in rows[i0 + 2:]:
        if len(r) < len(h):
            continue
        print("%-44s %9.1f %9.1f %9.1f %10.3g %7.1f %7.1f %5.0f %9s" % (
            short(r[col["Kernel Name"]])[:44], get(r, "gpu__time_duration.sum", to_us), get(r, "dram__bytes_read.sum", to_bytes) / 1e6,
            get(r, "dram__bytes_write.sum", to_bytes) / 1e6, get(r, "smsp__inst_executed.sum"),
            get(r, "smsp__issue_active.avg.pct_of_peak_sustained_active"), get(r, "sm__warps_active.avg.pct_of_peak_sustained_active"),
            get(r, "launch__registers_per_thread"), r[col["launch__grid_size"]] if "launch__grid_size" in col else ""))


def launches(path):
    rows = [r for r in csv.reader(open(path)) if r]
    i0 = next(i for i, r in enumerate(rows) if "Kernel Name" in r)
    h = rows[i0]
    kn, mv, mu = h.index("Kernel Name"), h.index("Metric Value"), h.index("Metric Unit")
    tot = OrderedDict()
    for r in rows[i0 + 1:]:
        if len(r) <= mv:
            continue
        k = short(r[kn])
        t = to_us(r[mv], r[mu])
        a = tot.setdefault(k, [0, 0.0])
        a[0] += 1; a[1] += t
    total = sum(a[1] for a in tot.values())
    print("%-44s %8s %12s %12s %7s" % ("kernel", "launches", "total ms", "avg us", "share"))
    for k, (n, t) in sorted(tot.items(), key=lambda kv: -kv[1][1]):
        print("%-44s %8d %12.3f %12.1f %6.1f%%" % (k[:44], n, t / 1e3, t / n, 100 * t / total))
    print("%-44s %8d %12.3f" % ("all", sum(a[0] for a in tot.values()), total / 1e3))


if __name__ == "__main__":
    (raw if sys.argv[1] == "raw" else launches)(sys.argv[2])
