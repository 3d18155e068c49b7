"""Per-source-line stall samples of one kernel of an .ncu-rep (ncu --page source --print-source cuda), top lines.
argv: report  kernel-id (e.g. ::fused_loop_kernel:2)  [top N]"""
import csv, io, subprocess, sys
rep, kid = sys.argv[1], sys.argv[2]
top = int(sys.argv[3]) if len(sys.argv) > 3 else 40
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass", "--kernel-id", kid],
                     capture_output=True, text=True).stdout
rows, cur_file, hdr = [], None, None
for r in csv.reader(io.StringIO(out)):
    if not r:
        continue
    if r[0] in ("File Name", "File Path"):
        cur_file = r[1].split("/")[-1]; hdr = None; continue
    if r[0] == "Function Name":
        continue
    if r[0] == "Line No":
        hdr = r; continue
    if hdr and cur_file and r[0].isdigit():
        if len(r) > 2 and r[2] != "-":
            continue                      # a SASS row under the line: the line's own row carries the totals
        d = {}
        for k, v in zip(hdr, r):
            d.setdefault(k, v)
        try:
            s = int(d.get("# Samples", "0") or 0)
        except ValueError:
            s = 0
        rows.append((s, cur_file, int(r[0]), d.get("Source", "")[:110], d))
tot = sum(r[0] for r in rows) or 1
print(f"# {kid}: {tot} samples")
for s, f, ln, src, d in sorted(rows, key=lambda t: -t[0])[:top]:
    stalls = {k[6:]: int(v) for k, v in d.items() if k.startswith("stall_") and "Not Issued" not in k and v.isdigit() and int(v) > 0}
    best = sorted(stalls.items(), key=lambda t: -t[1])[:3]
    print(f"{100.0*s/tot:5.1f}%  {f}:{ln:<4d} {src.strip()[:90]:90s} {best}")
