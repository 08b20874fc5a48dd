"""Dev: text summary of an .ncu-rep (read on the CPU box): per-kernel key metrics + the instructions that dominate
shared-memory wavefronts / stall samples.  usage: ncu_summary.py report.ncu-rep > profiles/xxx.txt"""
import csv
import io
import subprocess
import sys

KEYS = [
    "gpu__time_duration.sum", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
    "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_tensor",
    "smsp__issue_active.avg.pct", "sm__warps_active.avg.pct_of_peak_sustained_active",
    "dram__bytes_read.sum", "dram__bytes_write.sum", "dram__throughput.avg.pct_of_peak_sustained_elapsed",
    "lts__t_sector_hit_rate.pct", "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum",
    "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum.pct_of_peak_sustained_elapsed",
    "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "smsp__inst_executed.sum",
    "sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active", "sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active",
    "launch__registers_per_thread", "launch__grid_size", "launch__block_size", "launch__shared_mem_per_block_dynamic",
    "smsp__average_warp_latency_issue_stalled", "smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_math_pipe_throttle_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_not_selected_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_sleeping_per_issue_active.ratio",
]


def run(args):
    return subprocess.run(["ncu", "-i", rep] + args, capture_output=True, text=True).stdout


rep = sys.argv[1]
raw = list(csv.reader(io.StringIO(run(["--page", "raw", "--csv"]))))
hdr, units, rows = raw[0], raw[1], raw[2:]
ix = {h: i for i, h in enumerate(hdr)}
print(f"# ncu summary of {rep.split('/')[-1]} (ncu --set full --clock-control none; read with ncu -i ... --page raw/source --csv)")
for r in rows:
    name = r[ix["Kernel Name"]]
    print(f"\n== {name[:110]}")
    for k in KEYS:
        if k in ix and r[ix[k]] != "":
            print(f"   {k:90s} {r[ix[k]]:>18s} {units[ix[k]]}")
    rd, wr = ix.get("dram__bytes_read.sum"), ix.get("dram__bytes_write.sum")

# source page of the first kernel: top instructions by stall samples and by shared-memory wavefronts
txt = run(["--page", "source", "--csv"])
blocks = txt.split('"Kernel Name",')
for blk in blocks[1:]:
    lines = list(csv.reader(io.StringIO('"Kernel Name",' + blk)))
    kname = lines[0][1][:100]
    h = lines[1]
    hx = {c: i for i, c in enumerate(h)}
    data = [l for l in lines[2:] if len(l) == len(h)]

    def f(l, k):
        try:
            return float(l[hx[k]])
        except (ValueError, KeyError):
            return 0.0
    tot = sum(f(l, "# Samples") for l in data) or 1.0
    wf = sum(f(l, "L1 Wavefronts Shared") for l in data)
    print(f"\n== source counters: {kname}")
    print(f"   instructions executed (warp level) {sum(f(l, 'Instructions Executed') for l in data):.4g}; "
          f"shared-memory wavefronts attributed to LDS/STS/LDGSTS/ATOMS {wf:.4g} (excessive {sum(f(l, 'L1 Wavefronts Shared Excessive') for l in data):.4g})")
    print("   top 12 by stall samples:")
    for l in sorted(data, key=lambda l: -f(l, "# Samples"))[:12]:
        print(f"     {100 * f(l, '# Samples') / tot:5.1f} %  {l[hx['Source']].strip()[:90]}")
    print("   top 12 by shared-memory wavefronts:")
    for l in sorted(data, key=lambda l: -f(l, "L1 Wavefronts Shared"))[:12]:
        if f(l, "L1 Wavefronts Shared") > 0:
            print(f"     {f(l, 'L1 Wavefronts Shared'):10.4g} (excess {f(l, 'L1 Wavefronts Shared Excessive'):.3g})  {l[hx['Source']].strip()[:90]}")
