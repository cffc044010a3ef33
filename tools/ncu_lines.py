"""Top source lines by warp-stall samples from an ncu report (cuda,sass correlated view)."""
import csv, subprocess, sys
rep, kernels = sys.argv[1], sys.argv[2:]
for k in kernels:
    out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--print-source", "cuda,sass", "--csv", "--kernel-name", k],
                         capture_output=True, text=True).stdout
    rows = list(csv.reader(out.splitlines()))
    data, fname, hdr = [], "", None
    for r in rows:
        if not r: continue
        if r[0] == "File Path": fname = r[1].split("/")[-1]; continue
        if r[0] == "Line No": hdr = r; continue
        if hdr is None or len(r) < len(hdr) - 5: continue
        if r[2] != "-": continue   # keep the per-source-line aggregate rows only
        try:
            n = int(r[hdr.index("# Samples")]); ins = int(r[hdr.index("Instructions Executed")])
        except Exception: continue
        data.append((n, ins, fname, r[0], r[1].strip()[:100]))
    tot = sum(d[0] for d in data) or 1
    print(f"===== {k}: {tot} samples")
    for n, ins, f, l, s in sorted(data, key=lambda d: -d[0])[:18]:
        print(f"{100*n/tot:5.1f}% inst={ins:>9d} {f}:{l:>4s} {s}")
