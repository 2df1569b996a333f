"""Summarise an ncu --page raw --csv dump into profiles/<tag>_ncu_summary.md.
usage: python tools/ncu_summary.py gpurun_out/prof_X.ncu-rep <tag> "<title>" ["<reading>"]"""
import csv, subprocess, sys, io
rep, tag, title = sys.argv[1], sys.argv[2], sys.argv[3]
reading = sys.argv[4] if len(sys.argv) > 4 else ""
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
H, units = rows[0], rows[1]
want = ['Kernel Name', 'gpu__time_duration.sum', 'dram__bytes_read.sum', 'dram__bytes_write.sum',
        'gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed', 'sm__throughput.avg.pct_of_peak_sustained_elapsed',
        'sm__warps_active.avg.pct_of_peak_sustained_active', 'launch__registers_per_thread', 'smsp__inst_executed.sum',
        'smsp__issue_active.avg.pct_of_peak_sustained_active', 'lts__t_sector_hit_rate.pct', 'l1tex__t_sector_hit_rate.pct',
        'smsp__thread_inst_executed_per_inst_executed.ratio', 'lts__t_bytes.sum', 'l1tex__t_bytes.sum',
        'launch__grid_size', 'launch__block_size', 'launch__shared_mem_per_block_static']
with open(f"profiles/{tag}_ncu_summary.md", "w") as f:
    f.write(f"# {tag} — {title}\n\n")
    for r in rows[2:]:
        f.write("```\n")
        for w in want:
            if w in H:
                i = H.index(w)
                f.write(f"{w:70s} {r[i]:>24s} {units[i]}\n")
        f.write("```\n\n")
    if reading:
        f.write(reading + "\n")
# per-launch DRAM traffic of each kernel (dram__bytes_read.sum + dram__bytes_write.sum): bench.py reports it as roofline.traffic
import json, re
def _bytes(v, u):
    x = float(v.replace(",", ""))
    return x * {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}.get(u, 1)
traffic = {}
ir, iw, ik = H.index('dram__bytes_read.sum'), H.index('dram__bytes_write.sum'), H.index('Kernel Name')
for r in rows[2:]:
    name = re.sub(r"\(.*", "", r[ik]).replace("void ", "").strip()
    traffic.setdefault(name, []).append(_bytes(r[ir], units[ir]) + _bytes(r[iw], units[iw]))
def _git(*a):
    return subprocess.run(["git", *a], capture_output=True, text=True).stdout.strip()
commit = _git("rev-parse", "--short", "HEAD") + ("+dirty" if _git("status", "--porcelain", "--", "vk_renderer_b200/csrc", "include") else "")
with open(f"profiles/{tag}_traffic.json", "w") as f:
    json.dump({"kernel_source_commit": commit, "source": f"ncu --set full capture {tag} (bench.py --steps 2 --warmup 3, config 4); bytes per launch, mean over captured launches",
               "dram_bytes_per_launch": {k: sum(v) / len(v) for k, v in traffic.items()}}, f, indent=1)
print(open(f"profiles/{tag}_ncu_summary.md").read())
