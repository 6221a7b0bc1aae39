"""Turn the ncu outputs a gpurun call left in gpurun_out/ into the tracked summaries under profiles/.

    python scripts/summarise_ncu.py list   gpurun_out/launches.csv              profiles/r1_ncu_launch_list_final.txt
    python scripts/summarise_ncu.py full   gpurun_out/r1_final_gemm_T{16,512}.raw.csv   profiles/r1_ncu_full_summary_final.txt

`list`: the `--metrics gpu__time_duration.sum` launch list, aggregated per (kernel, grid, block).
`full`: selected metrics of `ncu -i X.ncu-rep --page raw --csv` dumps, one block per captured launch.
"""
import collections
import csv
import re
import sys

KEYS = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
        "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_tensor",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "launch__registers_per_thread", "launch__grid_size",
        "launch__block_size", "launch__cluster_size", "launch__shared_mem_per_block_dynamic", "sm__cycles_elapsed.avg",
        "smsp__inst_executed.sum", "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum", "lts__t_sector_hit_rate.pct",
        "smsp__issue_active.avg.pct_of_peak_sustained_active"]


def launch_list(src, dst):
    with open(src) as f:
        rows = list(csv.DictReader([l for l in f if l.startswith('"')]))
    agg = collections.OrderedDict()
    for x in rows:
        name = re.sub(r"\(.*", "", x["Kernel Name"])[:90]
        agg.setdefault((name, x["Grid Size"], x["Block Size"]), []).append(float(x["Metric Value"].replace(",", "")))
    out = ["# ncu --metrics gpu__time_duration.sum --clock-control none -c 400 python bench.py --steps 20 --warmup 3 --no-sweep --no-cpu",
           "# Per-launch times are cold-cache and serialised (no PDL overlap): compare SHARES of the step, not absolutes.",
           f"# unit: {rows[0]['Metric Unit']}",
           f"{'kernel':92s} {'grid':>14s} {'block':>12s} {'n':>4s} {'avg':>10s} {'min':>10s} {'max':>10s}"]
    for (n, g, b), v in agg.items():
        out.append(f"{n:92s} {g:>14s} {b:>12s} {len(v):4d} {sum(v) / len(v):10.1f} {min(v):10.1f} {max(v):10.1f}")
    open(dst, "w").write("\n".join(out) + "\n")
    print("\n".join(out))


def full(srcs, dst):
    out = ["# ncu --set full --clock-control none --import-source on -k regex:^kernel$ -s 3 -c 2, one GPU,",
           "# `python scripts/prof_one.py <T> 6` (4096->14336 I2_V_4).  Per launch.  Profiler timings are cold-cache, serialised and",
           "# without PDL overlap: bench.py's CUDA-event numbers are the ones to quote."]
    for src in srcs:
        rows = list(csv.reader(open(src)))
        hdr, units = rows[0], rows[1]
        out.append(f"## {src.split('/')[-1].replace('.raw.csv', '')}")
        for r in rows[2:]:
            d, u = dict(zip(hdr, r)), dict(zip(hdr, units))
            out.append(f"kernel: {d['Kernel Name'][:100]}  grid {d.get('Grid Size')} block {d.get('Block Size')}")
            for k in hdr:
                if any(k == kk or k.startswith(kk) for kk in KEYS):
                    out.append(f"  {k} = {d[k]} {u[k]}")
    open(dst, "w").write("\n".join(out) + "\n")
    print("\n".join(out[:60]))


if __name__ == "__main__":
    if sys.argv[1] == "list":
        launch_list(sys.argv[2], sys.argv[3])
    else:
        full(sys.argv[2:-1], sys.argv[-1])
