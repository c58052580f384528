"""Print headline metrics and the top stall sites of an .ncu-rep (reads it with `ncu -i`)."""
import csv, subprocess, sys, io
rep = sys.argv[1]
ntop = int(sys.argv[2]) if len(sys.argv) > 2 else 25
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr = rows[0]
keys = ['gpu__time_duration.sum', 'dram__bytes_read.sum', 'dram__bytes_write.sum',
        'gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed', 'sm__throughput.avg.pct_of_peak_sustained_elapsed',
        'sm__warps_active.avg.pct_of_peak_sustained_active', 'launch__registers_per_thread',
        'smsp__issue_active.avg.pct_of_peak_sustained_active', 'sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed',
        'l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum', 'smsp__inst_executed.sum', 'sm__cycles_elapsed.max',
        'sm__inst_executed_pipe_xu', 'sm__pipe_xu_cycles_active', 'sm__pipe_fma_cycles_active.avg.pct', 'sm__pipe_alu_cycles_active.avg.pct',
        'sm__inst_executed_pipe_fma', 'sm__inst_executed_pipe_alu', 'sm__inst_executed_pipe_lsu']
names = [r[hdr.index('Kernel Name')] for r in rows[2:]] if 'Kernel Name' in hdr else []
print(names)
for i, h in enumerate(hdr):
    if any(h == k or (h.startswith(k) and ('pct' in h or h.endswith('.sum'))) for k in keys):
        print(h, [r[i] for r in rows[1:]])
src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(src)))
hdr = None; out = []; kern = None
for r in rows:
    if r and r[0] == 'Kernel Name':
        if out: 
            pass
        kern = r[1]; continue
    if len(r) > 2 and r[0] == 'Address':
        hdr = r; si = hdr.index('# Samples'); continue
    if hdr and len(r) > si and r[0].startswith('0x'):
        try: out.append((kern, int(r[si]), r[1].strip(), r))
        except ValueError: pass
stall_cols = [i for i, h in enumerate(hdr) if h.startswith('stall_') and 'Not Issued' not in h]
for k in sorted(set(o[0] for o in out)):
    sub = [o for o in out if o[0] == k]
    tot = sum(o[1] for o in sub)
    print("==", k, "samples", tot)
    agg = {}
    for _, s, _, r in sub:
        for i in stall_cols:
            v = int(r[i]) if r[i].isdigit() else 0
            agg[hdr[i]] = agg.get(hdr[i], 0) + v
    print(sorted(agg.items(), key=lambda kv: -kv[1])[:8])
    for _, s, srcl, r in sorted(sub, key=lambda o: -o[1])[:ntop]:
        st = sorted(((int(r[i]) if r[i].isdigit() else 0, hdr[i]) for i in stall_cols), reverse=True)[:2]
        print(s, srcl[:80], st)
