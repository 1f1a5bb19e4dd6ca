"""Turn ncu outputs brought back in gpurun_out/ into the small text summaries committed under profiles/.

    python tools/ncu_summary.py launches gpurun_out/launches.csv            > profiles/rNN_launches.txt
    python tools/ncu_summary.py raw      gpurun_out/prof.ncu-rep [regex]    > profiles/rNN_kernel_raw.txt
"""
import collections
import csv
import io
import re
import subprocess
import sys

RAW_KEYS = (
    "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "gpu__dram_throughput.avg.pct",
    "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed", "sm__warps_active.avg.pct",
    "launch__registers_per_thread", "launch__grid_size", "launch__block_size", "sm__throughput.avg.pct",
    "l1tex__t_sector_hit_rate.pct", "lts__t_sector_hit_rate.pct", "smsp__inst_executed.sum",
    "smsp__issue_active.avg.pct", "sm__cycles_elapsed.max", "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum",
    "l1tex__data_pipe_lsu_wavefronts.sum", "lts__t_bytes.sum", "l1tex__t_bytes.sum",
    "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_elapsed", "launch__shared_mem_per_block_dynamic",
    "sm__inst_executed_pipe_fma.avg.pct", "sm__inst_executed_pipe_alu.avg.pct", "sm__inst_executed_pipe_fmaheavy",
    "l1tex__lsu_writeback_active", "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum",
)


def launches(path):
    lines = [l for l in open(path) if not l.startswith("==")]
    agg = collections.OrderedDict()
    tot = 0.0
    n = 0
    for row in csv.DictReader(lines):
        if row.get("Metric Name") != "gpu__time_duration.sum":
            continue
        v = float(row["Metric Value"].replace(",", ""))
        unit = row["Metric Unit"]
        v = v / 1e3 if unit in ("ns", "nsecond") else (v * 1e3 if unit in ("ms", "msecond") else v)
        name = row["Kernel Name"]
        m = re.search(r"(?:<unnamed>::|cpoly::)+(\w+)", name)   # our kernels: bare name without namespace / template args
        name = m.group(1) if m else re.sub(r"<.*", "", re.sub(r"\(.*", "", name))[:90]
        e = agg.setdefault(name, [0, 0.0])
        e[0] += 1
        e[1] += v
        tot += v
        n += 1
    print(f"# {path}: {n} launches, {tot:.1f} us summed device time (ncu: cold cache, serialised - compare SHARES)")
    for k, (c, t) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        print(f"{t:10.1f} us {100 * t / tot:5.1f}%  n={c:4d}  {k}")


def raw(path, pattern=None):
    out = subprocess.run(["ncu", "-i", path, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    hdr, units = rows[0], rows[1]
    ki = hdr.index("Kernel Name")
    for r in rows[2:]:
        if pattern and not re.search(pattern, r[ki]):
            continue
        print(f"## {r[ki][:120]}")
        for i, c in enumerate(hdr):
            if any(c.startswith(k) for k in RAW_KEYS):
                print(f"{c:90s} {r[i]:>16s} {units[i]}")


if __name__ == "__main__":
    if sys.argv[1] == "launches":
        launches(sys.argv[2])
    else:
        raw(sys.argv[2], sys.argv[3] if len(sys.argv) > 3 else None)
