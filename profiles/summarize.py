"""Turn ncu outputs brought back in gpurun_out/ into the text summaries committed under profiles/.

  python profiles/summarize.py launches gpurun_out/launches.csv          # per-kernel share of a step
  python profiles/summarize.py kernel   gpurun_out/prof.ncu-rep          # key metrics of a --set full capture
"""
import collections
import csv
import re
import subprocess
import sys

KEYS = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "dram__bytes_read.sum.per_second",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "lts__throughput.avg.pct_of_peak_sustained_elapsed",
        "lts__t_sectors_srcunit_tex_op_read.sum", "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed",
        "sm__throughput.avg.pct_of_peak_sustained_elapsed", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "smsp__inst_executed.sum", "launch__registers_per_thread", "launch__grid_size", "launch__block_size",
        "sm__cycles_elapsed.max", "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum"]


def launches(path):
    lines = [l for l in open(path) if not l.startswith("==")]
    rows = list(csv.DictReader(lines))
    agg = collections.OrderedDict()
    for x in rows:
        n = re.sub(r"\(.*", "", x["Kernel Name"])[:60]
        a = agg.setdefault(n, [0, 0.0])
        a[0] += 1
        a[1] += float(x["Metric Value"].replace(",", "")) / 1000.0
    tot = sum(v[1] for v in agg.values())
    print(f"| kernel | launches | total us | avg us | share |\n|---|---|---|---|---|")
    for n, (c, t) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        print(f"| `{n}` | {c} | {t:.1f} | {t / c:.1f} | {100 * t / tot:.1f}% |")


def kernel(path):
    out = subprocess.run(["ncu", "-i", path, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    r = list(csv.reader(out.splitlines()))
    hdr, units = r[0], r[1]
    for vals in r[2:]:
        name = vals[hdr.index("Kernel Name")] if "Kernel Name" in hdr else "?"
        print(f"### {name[:100]}\n")
        print("| metric | unit | value |\n|---|---|---|")
        for h, u, v in zip(hdr, units, vals):
            if h in KEYS:
                print(f"| {h} | {u} | {v} |")
        print()


if __name__ == "__main__":
    {"launches": launches, "kernel": kernel}[sys.argv[1]](sys.argv[2])
