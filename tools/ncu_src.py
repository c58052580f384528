"""Per-region view of an ncu source page: python tools/ncu_src.py REPORT KERNEL_REGEX [min_samples]
Prints the total instruction / sample counts, the stall reasons, and every SASS line with at least min_samples samples."""
import csv
import subprocess
import sys


def main():
    rep, rx = sys.argv[1], sys.argv[2]
    thr = int(sys.argv[3]) if len(sys.argv) > 3 else 40
    out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--kernel-name", "regex:" + rx, "--print-source", "sass"],
                         capture_output=True, text=True).stdout
    rows = list(csv.reader(out.splitlines()))
    hdr = next(r for r in rows if r and r[0] == "Address")
    ai, ie, si = hdr.index("Source"), hdr.index("Instructions Executed"), hdr.index("# Samples")
    stall = [i for i, h in enumerate(hdr) if h.startswith("stall_") and "Not Issued" not in h]
    body, started = [], False
    for r in rows:
        if r and r[0] == "Address":
            if started:
                break
            started = True
            continue
        if started and len(r) > ie and r[0].startswith("0x"):
            body.append(r)
    tot = sum(int(r[ie] or 0) for r in body)
    tots = sum(int(r[si] or 0) for r in body)
    print(f"{len(body)} lines, {tot} warp instructions, {tots} samples")
    agg = {}
    for r in body:
        for i in stall:
            v = int(r[i] or 0)
            if v:
                agg[hdr[i][6:]] = agg.get(hdr[i][6:], 0) + v
    print(" ".join(f"{k}:{v}" for k, v in sorted(agg.items(), key=lambda t: -t[1])))
    for n, r in enumerate(body):
        s = int(r[si] or 0)
        if s >= thr:
            why = sorted(((int(r[i] or 0), hdr[i][6:]) for i in stall), reverse=True)[:2]
            print(f"{n:5d} {int(r[ie] or 0) / tot * 100:5.2f}% inst {s:4d} samp  {why[0][1]}:{why[0][0]} {why[1][1]}:{why[1][0]}  {r[ai].strip()[:80]}")


if __name__ == "__main__":
    main()
