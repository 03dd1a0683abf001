"""Summarise ncu output into the tracked files under profiles/.

  python tools/ncu_summary.py launches <launches.csv> <out.md> "<title>"     per-kernel launch counts / time / share
  python tools/ncu_summary.py full <report.ncu-rep> <out.json>               key metrics per captured launch
The launch list comes from `ncu --metrics gpu__time_duration.sum --clock-control none --csv`; its per-launch times are
cold-cache and serialised, so only the SHARES are comparable with a plain run (B200_PROFILING.md).
"""
import collections
import csv
import io
import json
import re
import subprocess
import sys

KEYS = ["launch__grid_size", "launch__block_size", "launch__registers_per_thread", "gpu__time_duration.sum",
        "dram__bytes_read.sum", "dram__bytes_write.sum", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
        "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
        "sm__inst_executed_pipe_tensor.sum", "lts__t_sector_hit_rate.pct", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "smsp__cycles_active.avg", "sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active"]


def short(name):
    name = re.sub(r"\(.*", "", name)
    return name.replace("void ", "").replace("vlb::", "").strip()


def launches(path, out, title):
    text = open(path, errors="replace").read()
    start = text.find('"ID"')
    rows = list(csv.DictReader(io.StringIO(text[start:])))
    agg = collections.OrderedDict()
    for r in rows:
        if r.get("Metric Name") != "gpu__time_duration.sum":
            continue
        v = float(r["Metric Value"].replace(",", ""))
        unit = r.get("Metric Unit", "ns")
        v *= {"ns": 1e-6, "us": 1e-3, "ms": 1.0, "s": 1e3}.get(unit, 1e-6)
        k = short(r["Kernel Name"])
        a = agg.setdefault(k, [0, 0.0])
        a[0] += 1
        a[1] += v
    tot = sum(a[1] for a in agg.values())
    n = sum(a[0] for a in agg.values())
    with open(out, "w") as f:
        f.write("# %s\n\n" % title)
        f.write("`ncu --profile-from-start off --metrics gpu__time_duration.sum --clock-control none` (cold-cache, serialised: "
                "compare shares).\nSum of kernel times %.2f ms over %d launches.\n\n| kernel | launches | ms | share |\n|---|---|---|---|\n" % (tot, n))
        for k, a in sorted(agg.items(), key=lambda kv: -kv[1][1]):
            f.write("| `%s` | %d | %.3f | %.1f%% |\n" % (k, a[0], a[1], 100 * a[1] / tot))
    print("wrote", out, "total %.2f ms, %d launches" % (tot, n))


def full(rep, out):
    txt = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(txt)))
    hdr, units = rows[0], rows[1]
    res = []
    for r in rows[2:]:
        d = {"Kernel Name": short(r[hdr.index("Kernel Name")])}
        for k in KEYS:
            if k in hdr:
                i = hdr.index(k)
                d[k] = "%s %s" % (r[i], units[i])
        res.append(d)
    json.dump({"report": rep, "launches": res}, open(out, "w"), indent=1)
    print("wrote", out, len(res), "launches")


if __name__ == "__main__":
    if sys.argv[1] == "launches":
        launches(sys.argv[2], sys.argv[3], sys.argv[4] if len(sys.argv) > 4 else "ncu launch list")
    else:
        full(sys.argv[2], sys.argv[3])
