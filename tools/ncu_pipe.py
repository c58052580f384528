"""L1 data-pipe budget per kernel from an ncu report: LSU wavefronts (local/global and shared) + tensor-core operand
wavefronts against the elapsed cycles, beside issue, XU, DRAM.   usage: ncu_pipe.py REPORT.ncu-rep"""
import csv, io, subprocess, sys
raw = subprocess.run(["ncu", "-i", sys.argv[1], "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr = rows[0]
sms = 148
seen = set()
print("%-30s %7s %8s %8s %8s %8s %8s %6s %6s %6s %6s" % ("kernel", "us", "cycles", "lsu", "lgds", "shared", "tc", "pipe", "issue", "xu%", "dram%"))
for r in rows[2:]:
    d = dict(zip(hdr, r))
    n = d["Kernel Name"].replace("void ", "").replace("dlwp::", "").replace("<unnamed>::", "")[:30]
    if n in seen and "--all" not in sys.argv:
        continue
    seen.add(n)
    def f(k):
        try:
            return float(d.get(k, "").replace(",", ""))
        except ValueError:      # "", "n/a", "no data"
            return 0.0
    cyc = f("sm__cycles_elapsed.max")
    lsu = f("SM_A.TriageCompute.l1tex__data_pipe_lsu_wavefronts.avg")
    lg = f("SM_A.TriageCompute.l1tex__data_pipe_lsu_wavefronts_mem_lgds.avg")
    sh = f("SM_A.TriageCompute.l1tex__data_pipe_lsu_wavefronts_mem_shared.avg")
    tc = f("l1tex__data_pipe_tc_wavefronts_mem_shared.sum") / sms
    print("%-30s %7.1f %8.0f %8.0f %8.0f %8.0f %8.0f %6.2f %6.1f %6.1f %6.1f" % (
        n, f("gpu__time_duration.sum"), cyc, lsu, lg, sh, tc, (lsu + tc) / cyc if cyc else 0,
        f("smsp__issue_active.avg.pct_of_peak_sustained_active"), f("sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active"),
        f("gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed")))
