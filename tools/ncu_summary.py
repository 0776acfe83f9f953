"""Summarise an .ncu-rep (raw + source pages) into text: python tools/ncu_summary.py rep [ntop]"""
import csv
import io
import subprocess
import sys

rep = sys.argv[1]
ntop = int(sys.argv[2]) if len(sys.argv) > 2 else 40
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr, units, data = rows[0], rows[1], rows[2:]
keys = ['Kernel Name', 'gpu__time_duration.sum', 'dram__bytes_read.sum', 'dram__bytes_write.sum',
        'gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed',
        'sm__warps_active.avg.pct_of_peak_sustained_active', 'launch__registers_per_thread',
        'launch__grid_size', 'launch__block_size', 'launch__shared_mem_per_block_dynamic',
        'smsp__inst_executed.sum', 'sm__cycles_elapsed.avg',
        'smsp__issue_active.avg.pct_of_peak_sustained_active',
        'sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active',
        'l1tex__data_pipe_lsu_wavefronts_mem_shared.sum',
        'l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum',
        'l1tex__t_sector_hit_rate.pct', 'lts__t_sector_hit_rate.pct',
        'lts__t_sectors_srcunit_tex_op_read.sum', 'lts__t_bytes.sum',
        'l1tex__t_sectors_pipe_lsu_mem_global_op_ld.sum',
        'sm__inst_executed_pipe_lsu.sum', 'smsp__cycles_active.avg']
for k in keys:
    if k in hdr:
        i = hdr.index(k)
        print(f"{k} [{units[i]}]: {[r[i] for r in data]}")
src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(src)))
hi = [i for i, r in enumerate(rows) if r and r[0] == 'Address']
if not hi:
    sys.exit()
hdr = rows[hi[0]]
end = hi[1] - 1 if len(hi) > 1 else len(rows)
data = rows[hi[0] + 1:end]
iS, isrc, iex = hdr.index('# Samples'), hdr.index('Source'), hdr.index('Instructions Executed')
stalls = [h for h in hdr if h.startswith('stall_') and 'Not Issued' not in h]
tot = {s: 0 for s in stalls}
total = 0
for r in data:
    if not r[iS].isdigit():
        continue
    total += int(r[iS])
    for s in stalls:
        v = r[hdr.index(s)]
        if v.isdigit():
            tot[s] += int(v)
print('total samples', total, 'instructions', len(data))
for s, v in sorted(tot.items(), key=lambda x: -x[1])[:8]:
    print(f"  {s}: {v} ({100 * v / max(total, 1):.1f}%)")
top = sorted([r for r in data if r[iS].isdigit()], key=lambda r: -int(r[iS]))[:ntop]
for r in top:
    d = {s: int(r[hdr.index(s)]) for s in stalls if r[hdr.index(s)].isdigit() and int(r[hdr.index(s)]) > 0}
    best = sorted(d.items(), key=lambda x: -x[1])[:2]
    print(r[iS], r[iex], r[isrc][:80], best)
