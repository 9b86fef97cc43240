"""Turns ncu reports / launch lists brought back in gpurun_out/ into the text summaries kept in profiles/."""
import collections
import csv
import subprocess
import sys

KEYS = [
    "gpu__time_duration.sum", "sm__cycles_elapsed.max", "launch__grid_size", "launch__block_size",
    "launch__registers_per_thread", "launch__shared_mem_per_block_dynamic", "launch__occupancy_limit_shared_mem",
    "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
    "smsp__issue_active.avg.pct_of_peak_sustained_active", "smsp__inst_executed.sum",
    "sm__warps_active.avg.pct_of_peak_sustained_active", "dram__bytes_read.sum", "dram__bytes_write.sum",
    "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "lts__t_bytes.sum",
    "lts__throughput.avg.pct_of_peak_sustained_elapsed", "l1tex__throughput.avg.pct_of_peak_sustained_elapsed",
    "lts__t_sector_hit_rate.pct", "l1tex__t_sector_hit_rate.pct",
    "smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_wait_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_branch_resolving_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_lg_throttle_per_issue_active.ratio",
    "l1tex__data_pipe_lsu_wavefronts.sum.pct_of_peak_sustained_elapsed",
    "l1tex__data_pipe_tc_wavefronts_mem_shared.sum.pct_of_peak_sustained_elapsed",
    "l1tex__throughput.avg.pct_of_peak_sustained_active", "l1tex__m_xbar2l1tex_read_bytes.sum",
    "dram__throughput.avg.pct_of_peak_sustained_elapsed",
]


def raw(rep):
    out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(out.splitlines()))
    hdr, units = rows[0], rows[1]
    res = []
    for vals in rows[2:]:
        d = {h: (vals[i], units[i]) for i, h in enumerate(hdr)}
        res.append(d)
    return res


def report(rep):
    for d in raw(rep):
        print("kernel:", d["Kernel Name"][0][:110])
        for k in KEYS:
            if k in d:
                print(f"  {k:90s} {d[k][0]:>16s} {d[k][1]}")
        print()


def launches(path):
    rows = [r for r in csv.reader(open(path)) if len(r) > 10 and r[0].isdigit()]
    tot = collections.defaultdict(float)
    cnt = collections.Counter()
    for r in rows:
        name = r[4].split("(")[0][:70]
        if "gate_kernel" in name:   # bench.py's start gate (under ncu, which serialises launches, it runs into its time limit)
            continue
        tot[name] += float(r[-1])
        cnt[name] += 1
    s = sum(tot.values())
    print(f"{'kernel':72s} {'launches':>8s} {'total us':>10s} {'avg us':>9s} {'share':>6s}")
    for k, v in sorted(tot.items(), key=lambda kv: -kv[1]):
        print(f"{k:72s} {cnt[k]:8d} {v / 1e3:10.1f} {v / 1e3 / cnt[k]:9.1f} {100 * v / s:5.1f}%")


if __name__ == "__main__":
    if sys.argv[1] == "launches":
        launches(sys.argv[2])
    else:
        report(sys.argv[2] if len(sys.argv) > 2 else sys.argv[1])
