"""Per-launch table of one forward pass from an ncu --csv log (tools/gpu_r2_ncu.sh): duration, DRAM bytes read / written and
the tensor-pipe metrics collected, one row per kernel launch in launch order, plus totals and -- with --json -- the
roofline.traffic file bench.py reads (profiles/r2_dram_traffic.json, tagged with the kernel source id).
  python tools/ncu_layers.py gpurun_out/layers.csv [--ops gpurun_out/ops.txt] [--json profiles/r2_dram_traffic.json]"""
import argparse, collections, csv, hashlib, json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ap = argparse.ArgumentParser()
ap.add_argument("csv")
ap.add_argument("--ops", default="")
ap.add_argument("--json", default="")
a = ap.parse_args()
rows = [r for r in csv.reader(open(a.csv)) if len(r) > 14 and r[0].isdigit()]
per = collections.OrderedDict()
for r in rows:
    per.setdefault(int(r[0]), {"k": r[4]})[r[12]] = float(r[14].replace(",", "")) if r[14] not in ("", "n/a") else float("nan")
ops = []
if a.ops and os.path.exists(a.ops):
    ops = [l.split("  ", 3)[-1].strip()[:64] for l in open(a.ops) if l[:4].strip().isdigit()]
metrics = sorted({m for d in per.values() for m in d if m != "k"})
short = {m: m.replace("sm__", "").replace(".avg.pct_of_peak_sustained_elapsed", "%").replace("gpu__time_duration.sum", "ns")
             .replace("dram__bytes_", "dram_").replace(".sum", "") for m in metrics}
print("launch " + " ".join(f"{short[m]:>22s}" for m in metrics) + "  kernel / op")
tot = collections.Counter()
for n, (i, d) in enumerate(per.items()):
    name = d["k"].split("(")[0].split("::")[-1]
    print(f"{n:6d} " + " ".join(f"{d.get(m, float('nan')):22.1f}" for m in metrics) + f"  {name} | {ops[n] if n < len(ops) else ''}")
    for m in metrics:
        if d.get(m) == d.get(m):
            tot[m] += d.get(m, 0)
print(f"{'TOTAL':6s} " + " ".join(f"{tot[m]:22.1f}" for m in metrics))
t = "gpu__time_duration.sum"
for m in metrics:
    if m.endswith("pct_of_peak_sustained_elapsed") and t in tot:
        w = sum(d.get(m, 0) * d.get(t, 0) for d in per.values() if d.get(m) == d.get(m)) / tot[t]
        print(f"time-weighted {m}: {w:.2f} %")
if a.json:
    h = hashlib.sha256()
    for name in ("conv_igemm.cu", "conv_igemm.cuh", "ptx.cuh", "aux_kernels.cu", "api.cu"):
        h.update(open(os.path.join(ROOT, "aquaculture_b200", "csrc", name), "rb").read())
    conv = [d for d in per.values() if "conv_igemm" in d["k"]]
    out = dict(source=a.csv, kernel_source_id=h.hexdigest()[:16], launches=len(per), conv_stack_launches=len(conv),
               conv_stack_dram_bytes=sum(d.get("dram__bytes_read.sum", 0) + d.get("dram__bytes_write.sum", 0) for d in conv),
               conv_stack_dram_read=sum(d.get("dram__bytes_read.sum", 0) for d in conv),
               conv_stack_dram_write=sum(d.get("dram__bytes_write.sum", 0) for d in conv),
               all_launches_ns=tot.get(t, 0))
    json.dump(out, open(a.json, "w"), indent=1)
    print(json.dumps(out))
