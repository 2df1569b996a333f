"""Per-CUDA-source-line instruction and stall-sample totals from an ncu report (needs -lineinfo).
usage: python tools/ncu_lines.py rep kernel_regex launch_skip [topN]"""
import csv, subprocess, sys, io
rep, kre, skip = sys.argv[1], sys.argv[2], sys.argv[3]
top = int(sys.argv[4]) if len(sys.argv) > 4 else 40
raw = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass", "--kernel-name", f"regex:{kre}",
                      "--launch-skip", skip, "--launch-count", "1"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
out = []
cur_file = ""
H = None
for r in rows:
    if len(r) > 6 and r[0] == 'Line No':
        H = r; si = H.index('# Samples'); ie = H.index('Instructions Executed'); continue
    if H is None: continue
    if len(r) == 2 and r[0] == 'File Path' or (len(r) >= 2 and r[0].startswith('File')):
        cur_file = r[1] if len(r) > 1 else ''
        continue
    if len(r) > ie and r[0].isdigit() and r[2] == '-':
        try: out.append((int(r[ie]), int(r[si]), cur_file.split('/')[-1], int(r[0]), r[1].strip()))
        except ValueError: pass
ti = sum(o[0] for o in out); ts = sum(o[1] for o in out)
print(f"total inst {ti}  total samples {ts}")
for n, s, f, ln, src in sorted(out, reverse=True)[:top]:
    print(f"{100*n/ti:5.1f}% inst {100*s/max(ts,1):5.1f}% smp  {f}:{ln:4d}  {src[:110]}")
