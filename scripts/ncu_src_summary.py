"""ncu_src_summary.py report.ncu-rep [top]: summarise the source page of an ncu report (stall samples per SASS instruction).

Prints headline metrics, the most-sampled instructions with their top stall reasons, and every tcgen05 / TMA / mbarrier
instruction with its sample count (a coarse map of the warp roles)."""
import csv, subprocess, sys, io

rep = sys.argv[1]
top_n = int(sys.argv[2]) if len(sys.argv) > 2 else 40
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr, units, vals = rows[0], rows[1], rows[2]
keys = ['Kernel Name', 'gpu__time_duration.sum', 'dram__bytes_read.sum', 'dram__bytes_write.sum', 'lts__t_sector_hit_rate.pct',
        'sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active', 'l1tex__m_xbar2l1tex_read_bytes.sum',
        'lts__throughput.avg.pct_of_peak_sustained_elapsed', 'launch__registers_per_thread', 'smsp__inst_executed.sum',
        'gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed', 'l1tex__m_l1tex2xbar_write_bytes.sum',
        'launch__grid_size', 'sm__cycles_elapsed.avg', 'smsp__warp_issue_stalled_long_scoreboard_per_warp_active.pct']
for h, u, v in zip(hdr, units, vals):
    if h in keys:
        print(f"  {h} = {v} {u}")
src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(src)))
hdr = rows[1]
data = rows[2:]
ci = {h: i for i, h in enumerate(hdr)}
stall_cols = [h for h in hdr if h.startswith('stall_') and 'Not Issued' not in h]
tot = sum(int(r[ci['# Samples']]) for r in data)
print(f"  total warp-stall samples = {tot}, SASS instructions = {len(data)}")
def line(i):
    r = data[i]
    st = {h: int(r[ci[h]]) for h in stall_cols if int(r[ci[h]]) > 0}
    st = sorted(st.items(), key=lambda kv: -kv[1])[:3]
    return f"  {i:5d} {int(r[ci['# Samples']]):6d} {int(r[ci['Instructions Executed']]):8d}  {r[ci['Source']].strip()[:72]:72s} {st}"
print("# most-sampled instructions: row, samples, executions, SASS, top stall reasons")
for i in sorted(sorted(range(len(data)), key=lambda i: -int(data[i][ci['# Samples']]))[:top_n]):
    print(line(i))
print("# tcgen05 / TMA / mbarrier instructions")
for i, r in enumerate(data):
    s = r[ci['Source']]
    if any(k in s for k in ('UTCIMMA', 'UTCHMMA', 'UBLKCP', 'UTCBAR', 'STTM', 'LDTM', 'SYNCS', 'BAR.SYNC', 'ACQBULK', 'UCGABAR', 'MEMBAR')):
        print(line(i))
