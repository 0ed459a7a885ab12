"""Summarises ncu outputs brought back in gpurun_out/ into small text files for profiles/.

    python scripts/ncu_summary.py launches gpurun_out/launches.csv > profiles/rXX_launches.txt
    python scripts/ncu_summary.py full gpurun_out/prof.ncu-rep > profiles/rXX_kernel.txt
"""
import collections
import csv
import subprocess
import sys

KEYS = [
    'gpu__time_duration.sum', 'dram__bytes_read.sum', 'dram__bytes_write.sum',
    'gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed',
    'sm__throughput.avg.pct_of_peak_sustained_elapsed', 'smsp__issue_active.avg.pct_of_peak_sustained_active',
    'sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active',
    'sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active',
    'sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active',
    'sm__inst_executed_pipe_tensor.avg.pct_of_peak_sustained_active',
    'sm__warps_active.avg.pct_of_peak_sustained_active', 'launch__registers_per_thread',
    'launch__occupancy_limit_shared_mem', 'launch__occupancy_limit_registers', 'launch__grid_size',
    'launch__block_size', 'smsp__inst_executed.sum', 'l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum',
    'sm__cycles_elapsed.avg', 'lts__t_sector_hit_rate.pct', 'dram__cycles_active.avg.pct_of_peak_sustained_elapsed',
]


def launches(path):
    rows = [r for r in csv.reader(open(path, errors="replace")) if len(r) > 10]
    hdr = rows[0]
    ki, vi, ui = hdr.index('Kernel Name'), hdr.index('Metric Value'), hdr.index('Metric Unit')
    agg = collections.OrderedDict()
    unit = rows[1][ui]
    scale = {'ns': 1e-6, 'us': 1e-3, 'ms': 1.0, 's': 1e3}.get(unit, 1e-6)
    for r in rows[1:]:
        agg.setdefault(r[ki][:90], []).append(float(r[vi].replace(',', '')) * scale)
    tot = sum(sum(v) for v in agg.values())
    print(f"# per-kernel device time from {path} (ncu gpu__time_duration.sum, cold-cache, serialised)")
    print(f"# total {tot:.3f} ms over {sum(len(v) for v in agg.values())} launches")
    for k, v in sorted(agg.items(), key=lambda kv: -sum(kv[1])):
        print(f"{sum(v) / tot:6.3f}  n={len(v):4d}  total={sum(v):10.3f} ms  avg={sum(v) / len(v):9.4f} ms  {k}")


def full(path):
    out = subprocess.run(["ncu", "-i", path, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(out.splitlines()))
    h, units = rows[0], rows[1]
    print(f"# ncu --set full summary of {path}")
    for r in rows[2:]:
        name = r[h.index('Kernel Name')]
        print(f"## {name[:100]}")
        for k in KEYS:
            if k in h:
                i = h.index(k)
                print(f"{k:70s} {r[i]:>20s} {units[i]}")


if __name__ == "__main__":
    {"launches": launches, "full": full}[sys.argv[1]](sys.argv[2])
