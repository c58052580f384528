"""Top stall-sample instructions per kernel from `ncu -i REP --page source --csv` (first launch of each kernel name)."""
import csv, subprocess, sys, io
rep, pat = sys.argv[1], sys.argv[2]
n = int(sys.argv[3]) if len(sys.argv) > 3 else 12
raw = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--kernel-name", "regex:" + pat], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
secs, cur = [], None
for r in rows:
    if r and r[0] == "Kernel Name":
        cur = {"name": r[1], "rows": []}; secs.append(cur)
    elif cur is not None:
        cur["rows"].append(r)
seen = set()
for s in secs:
    if s["name"] in seen: continue
    seen.add(s["name"])
    hdr, data = s["rows"][0], [r for r in s["rows"][1:] if len(r) == len(s["rows"][0])]
    iS, iSrc = hdr.index("# Samples"), hdr.index("Source")
    tot = sum(int(r[iS]) for r in data)
    print(s["name"][:70], "samples", tot, "instrs", len(data))
    for i in sorted(sorted(range(len(data)), key=lambda i: -int(data[i][iS]))[:n]):
        print("  ", i, data[i][iS], data[i][iSrc][:100])
