"""Summarises an .ncu-rep: key raw metrics, stall-reason totals, instruction mix, top lines."""
import csv, subprocess, sys, collections, io
rep = sys.argv[1]
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
r = list(csv.reader(io.StringIO(raw)))
hdr, units, vals = r[0], r[1], r[2]
want = ['Kernel Name', 'gpu__time_duration.sum', 'dram__bytes_read.sum', 'dram__bytes_write.sum', 'launch__registers_per_thread', 'launch__grid_size',
        'launch__block_size', 'launch__shared_mem_per_block_dynamic', 'sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active',
        'sm__inst_executed_pipe_tensor_subpipe_dmma.avg.pct_of_peak_sustained_active', 'sm__warps_active.avg.pct_of_peak_sustained_active',
        'l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum', 'sm__throughput.avg.pct_of_peak_sustained_elapsed', 'lts__t_sector_hit_rate.pct',
        'smsp__inst_executed.sum', 'sm__cycles_elapsed.max', 'smsp__issue_active.avg.pct_of_peak_sustained_active', 'gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed']
for h, u, v in zip(hdr, units, vals):
    if h in want: print(f"{h} [{u}] = {v}")
src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(src)))
h = rows[1]; idx = {x: i for i, x in enumerate(h)}; data = rows[2:]
f = lambda x: float(x) if x not in ("", None) else 0.0
def ff(x):
    try: return float(x)
    except: return 0.0
tot = sum(ff(r[idx['# Samples']]) for r in data)
stalls = [x for x in h if x.startswith('stall_') and 'Not Issued' not in x]
agg = {s: sum(ff(r[idx[s]]) for r in data) for s in stalls}
print("samples", tot)
for k, v in sorted(agg.items(), key=lambda x: -x[1])[:8]: print(f"  {k:28s} {v / tot:.3f}")
mix = collections.Counter()
for r in data:
    op = r[idx['Source']].split()
    if not op: continue
    o = op[0] if not op[0].startswith('@') else (op[1] if len(op) > 1 else op[0])
    mix[o.split('.')[0]] += ff(r[idx['Instructions Executed']])
T = sum(mix.values())
print("instruction mix (warp-level):")
for k, v in mix.most_common(16): print(f"  {k:10s} {v:.3e} {v / T:.3f}")
print("top sampled instructions:")
for r in sorted(data, key=lambda r: -ff(r[idx['# Samples']]))[:int(sys.argv[2]) if len(sys.argv) > 2 else 12]:
    st = {s: ff(r[idx[s]]) for s in stalls}
    print(f"  {ff(r[idx['# Samples']]):8.0f} {max(st, key=st.get):22s} {r[idx['Source']][:80]}")
