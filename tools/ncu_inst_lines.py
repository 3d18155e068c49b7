"""Per-source-line INSTRUCTIONS EXECUTED (and samples) of one kernel of an .ncu-rep, top lines.
argv: report  kernel-id (e.g. ::vote_frame_kernel:4)  [top N]"""
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
            continue
        d = {}
        for k, v in zip(hdr, r):
            d.setdefault(k, v)
        def num(k):
            try:
                return int(d.get(k, "0") or 0)
            except ValueError:
                return 0
        rows.append((num("Instructions Executed"), num("# Samples"), cur_file, int(r[0]), d.get("Source", "")[:100]))
tot = sum(r[0] for r in rows) or 1
tots = sum(r[1] for r in rows) or 1
print(f"# {kid}: {tot} warp-instructions, {tots} samples")
for n, s, f, ln, src in sorted(rows, key=lambda t: -t[0])[:top]:
    print(f"{100.0*n/tot:5.1f}% inst {100.0*s/tots:5.1f}% smp  {f}:{ln:<4d} {src.strip()[:100]}")
