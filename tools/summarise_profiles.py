"""gpurun_out/{launches.csv, prof_full.ncu-rep, plain.log} -> profiles/<tag>/ (tracked summaries) + profiles/traffic.json."""
import collections, csv, json, os, shutil, subprocess, sys
tag = sys.argv[1] if len(sys.argv) > 1 else "r02_final"
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
src, dst = os.path.join(ROOT, "gpurun_out"), os.path.join(ROOT, "profiles", tag)
os.makedirs(dst, exist_ok=True)
shutil.copy(os.path.join(src, "launches.csv"), os.path.join(dst, "launches.csv"))
with open(os.path.join(src, "plain.log")) as f:
    line = [l for l in f if l.startswith("{")][-1]
with open(os.path.join(dst, "bench_line_same_command.json"), "w") as f:
    f.write(line)
# per-kernel share of the launch list (cold-cache, serialised: compare shares, not absolutes)
rows = list(csv.reader(l for l in open(os.path.join(dst, "launches.csv")) if not l.startswith("==")))
hdr = rows[0]; ki, vi, ui = hdr.index("Kernel Name"), hdr.index("Metric Value"), hdr.index("Metric Unit")
tot, cnt = collections.Counter(), collections.Counter()
for r in rows[1:]:
    if len(r) <= vi: continue
    v = float(r[vi].replace(",", "")); v = v / 1e6 if r[ui] == "ns" else (v / 1e3 if r[ui] == "us" else v)
    name = r[ki].split("(")[0].replace("void ", ""); tot[name] += v; cnt[name] += 1
s = sum(tot.values())
with open(os.path.join(dst, "launch_shares.txt"), "w") as f:
    f.write("ncu --metrics gpu__time_duration.sum --clock-control none, command: python bench.py --steps 2 --warmup 1 --no-cpu-baseline --no-e2e\n")
    f.write("(includes the one-off generator kernels gen_fill/gen_type/radix_hist + its 3 radix passes)\n")
    for k, v in tot.most_common():
        f.write("%-44s launches=%3d  total=%9.3f ms  share=%5.1f%%\n" % (k[:44], cnt[k], v, 100 * v / s))
summ = subprocess.run([sys.executable, os.path.join(ROOT, "tools", "ncu_summary.py"), os.path.join(src, "prof_full.ncu-rep")],
                      capture_output=True, text=True).stdout
open(os.path.join(dst, "ncu_full_summary.csv"), "w").write(summ)
for kern, fn in (("radix_pass", "pte_radix.cuh"), ("intern_pack", "pte_kernels.cuh"), ("diff_flag", "pte_kernels.cuh"),
                 ("assemble_bitmap", "pte_kernels.cuh")):
    out = subprocess.run([sys.executable, os.path.join(ROOT, "tools", "ncu_lines.py"), os.path.join(src, "prof_full.ncu-rep"), kern, fn],
                         capture_output=True, text=True).stdout
    open(os.path.join(dst, "ncu_lines_%s.txt" % kern), "w").write(out)
# traffic.json
raw = subprocess.run(["ncu", "-i", os.path.join(src, "prof_full.ncu-rep"), "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(raw.splitlines())); hdr, units = rows[0], rows[1]
ki, ri, wi = hdr.index("Kernel Name"), hdr.index("dram__bytes_read.sum"), hdr.index("dram__bytes_write.sum")
scale = {"Gbyte": 1e9, "Mbyte": 1e6, "Kbyte": 1e3, "byte": 1}
acc = {}
for r in rows[2:]:
    name = r[ki].split("(")[0].replace("void ", "").split("<")[0]
    acc.setdefault(name, []).append(float(r[ri]) * scale[units[ri]] + float(r[wi]) * scale[units[wi]])
json.dump({"source": "profiles/%s/ncu_full_summary.csv (ncu --set full --clock-control none, one capture per kernel of a timed step, "
                     "BASELINE config 3: 1B events / 10M pods)" % tag,
           "dram_bytes_per_launch": {k: sum(v) / len(v) for k, v in acc.items()}},
          open(os.path.join(ROOT, "profiles", "traffic.json"), "w"), indent=1)
print(open(os.path.join(dst, "launch_shares.txt")).read())
