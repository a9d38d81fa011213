"""Turn the raw ncu artefacts a GPU visit leaves in gpurun_out/ into the small, tracked summaries under profiles/.

    python tools/ncu_summarize.py launches gpurun_out/launches.csv  profiles/r01_launches.md
    python tools/ncu_summarize.py full     gpurun_out/prof_gemm.ncu-rep profiles/r01_gemm_full.md

`launches`: per-kernel launch count, summed gpu__time_duration and share of the captured command
(the per-launch times are cold-cache and serialised: the SHARE is what must agree with bench.py).
`full`: the handful of `--set full` metrics the roofline argument uses, per captured launch.
"""
import collections
import csv
import subprocess
import sys

FULL_KEYS = [
    "gpu__time_duration.sum",
    "launch__grid_size",
    "launch__block_size",
    "launch__cluster_size",
    "launch__registers_per_thread",
    "launch__shared_mem_per_block_dynamic",
    "sm__cycles_elapsed.max",
    "smsp__cycles_active.avg",
    "sm__throughput.avg.pct_of_peak_sustained_elapsed",
    "sm__pipe_tensor_cycles_active_realtime.avg.pct_of_peak_sustained_elapsed",
    "sm__pipe_tensor_subpipe_imma_cycles_active_realtime.avg",
    "sm__pipe_tensor_subpipe_hmma_cycles_active_realtime.avg",
    "sm__inst_executed_pipe_tmem.sum",
    "sm__inst_executed_pipe_uniform.sum",
    "dram__bytes_read.sum",
    "dram__bytes_write.sum",
    "dram__bytes_read.sum.per_second",
    "dram__bytes_write.sum.per_second",
    "dram__throughput.avg.pct_of_peak_sustained_elapsed",
    "lts__t_bytes.sum",
    "lts__t_sector_hit_rate.pct",
    "lts__throughput.avg.pct_of_peak_sustained_elapsed",
    "l1tex__m_xbar2l1tex_read_bytes.sum",
    "l1tex__m_xbar2l1tex_read_bytes.sum.per_second",
    "l1tex__m_xbar2l1tex_read_bytes.sum.pct_of_peak_sustained_elapsed",
    "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum",
    "smsp__warps_issue_stalled_long_scoreboard_per_warp_active.pct",
    "sm__warps_active.avg.pct_of_peak_sustained_active",
    "gpc__cycles_elapsed.avg.per_second",
    "sm__cycles_elapsed.avg.per_second",
]


def launches(src, dst):
    rows = [r for r in csv.reader(open(src)) if len(r) > 10]
    hdr = rows[0]
    ki, vi, gi, bi = hdr.index("Kernel Name"), hdr.index("Metric Value"), hdr.index("Grid Size"), hdr.index("Block Size")
    agg = collections.OrderedDict()
    for r in rows[1:]:
        name = r[ki]
        short = name.split("(")[0].replace("void ", "")[:100]
        key = (short, r[gi], r[bi])
        a = agg.setdefault(key, [0, 0.0])
        a[0] += 1
        a[1] += float(r[vi].replace(",", ""))
    tot = sum(a[1] for a in agg.values())
    with open(dst, "w") as f:
        f.write(f"# ncu launch list ({src})\n\n`ncu --metrics gpu__time_duration.sum --clock-control none` of the bench command; "
                "times are cold-cache and serialised, so read the SHARE column.\n\n")
        f.write("| kernel | grid | block | launches | total us | share |\n|---|---|---|---|---|---|\n")
        for (k, g, b), (n, t) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
            f.write(f"| `{k}` | {g} | {b} | {n} | {t / 1e3:.1f} | {100 * t / tot:.1f}% |\n")
        f.write(f"\nTotal captured GPU time: {tot / 1e3:.1f} us over {sum(a[0] for a in agg.values())} launches.\n")


def full(src, dst):
    raw = subprocess.run(["ncu", "-i", src, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(raw.splitlines()))
    hdr, units, data = rows[0], rows[1], rows[2:]
    with open(dst, "w") as f:
        f.write(f"# ncu --set full summary ({src})\n\n")
        ni = hdr.index("Kernel Name")
        f.write("Launches: " + "; ".join(f"[{i}] `{r[ni][:80]}`" for i, r in enumerate(data)) + "\n\n")
        f.write("| metric | unit | " + " | ".join(f"[{i}]" for i in range(len(data))) + " |\n")
        f.write("|---|---|" + "---|" * len(data) + "\n")
        for key in FULL_KEYS:
            for i, h in enumerate(hdr):
                if h == key or h.endswith("." + key):
                    f.write(f"| {h} | {units[i]} | " + " | ".join(r[i] for r in data) + " |\n")
                    break


if __name__ == "__main__":
    {"launches": launches, "full": full}[sys.argv[1]](sys.argv[2], sys.argv[3])
