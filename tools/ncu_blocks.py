"""Per-kernel instruction totals of an .ncu-rep, grouped into runs of SASS instructions that
executed the same number of times (= loop levels).  Usage: python tools/ncu_blocks.py rep [min_share] [kernel_index dump_lo dump_hi]"""
import csv, subprocess, sys
rep = sys.argv[1]
share = float(sys.argv[2]) if len(sys.argv) > 2 else 0.012
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass"], capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
starts = [i for i, r in enumerate(rows) if r and r[0] == "Kernel Name"]
seen = set()
for si, st in enumerate(starts):
    end = starts[si + 1] if si + 1 < len(starts) else len(rows)
    name = rows[st][1]
    if name in seen: continue
    seen.add(name)
    hi = [i for i in range(st, end) if rows[i] and rows[i][0] == "Address"][0]
    H = rows[hi]; ie = H.index("Instructions Executed"); isrc = H.index("Source")
    data = rows[hi + 1:end]
    blocks = []
    for k, r in enumerate(data):
        n = int(r[ie]) if len(r) > ie and r[ie].isdigit() else 0
        if blocks and blocks[-1][2] == n: blocks[-1][1] = k
        else: blocks.append([k, k, n])
    tot = sum((b - a + 1) * n for a, b, n in blocks)
    print("==== kernel", si, name[:70], "total %.1fM" % (tot / 1e6))
    for a, b, n in blocks:
        if (b - a + 1) * n > tot * share:
            print(f"{a:4d}-{b:4d} len {b-a+1:3d} count {n/1e6:7.2f}M total {(b-a+1)*n/1e6:7.1f}M   first: {data[a][isrc][:46]}")
    if len(sys.argv) > 5 and int(sys.argv[3]) == si:
        for k in range(int(sys.argv[4]), int(sys.argv[5])):
            r = data[k]; print(f"{k:4d} {int(r[ie])/1e6:6.2f}M {r[isrc][:90]}")
