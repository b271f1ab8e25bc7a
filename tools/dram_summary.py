"""Per-step DRAM traffic from the ncu pass tools/gpu_dram_pass.sh writes (gpu__time_duration + dram read/write bytes
per launch of `bench.py --steps 2 --warmup 3`): python tools/dram_summary.py gpurun_out/dram_<tag>.csv [out.json]
A step starts at each preprocess_kernel launch; the LAST complete step is summarised per kernel family.  The JSON
feeds bench.py's roofline.traffic (bytes per conv-stack launch set, i.e. per batch of 64 tiles)."""
import collections
import csv
import json
import sys

rows = [r for r in csv.reader(open(sys.argv[1])) if len(r) > 14 and r[0].isdigit()]
per = collections.OrderedDict()
for r in rows:
    per.setdefault(int(r[0]), {"k": r[4]})[r[12]] = float(r[14])
ids = sorted(per)
starts = [i for i in ids if "preprocess" in per[i]["k"]]
# steps: [starts[j], starts[j+1]); drop everything after the last preprocess unless it is a complete step
lens = [b - a for a, b in zip(starts, starts[1:]) if b - a > 10]  # back-to-back preprocess launches = its own timing loop
step_len = max(set(lens), key=lens.count)
cands = [a for a, b in zip(starts, starts[1:]) if b - a == step_len]
# the timed e2e passes come last in bench.py; take the first step after warm-up of the resident loop (3 warm-ups)
a = cands[min(3, len(cands) - 1)]
agg = collections.OrderedDict()
for i in range(a, a + step_len):
    d = per[i]
    k = d["k"].split("(")[0].split("::")[-1]
    g = agg.setdefault(k, dict(launches=0, us=0.0, read=0.0, write=0.0))
    g["launches"] += 1
    g["us"] += d.get("gpu__time_duration.sum", 0) / 1e3
    g["read"] += d.get("dram__bytes_read.sum", 0)
    g["write"] += d.get("dram__bytes_write.sum", 0)
tot_us = sum(g["us"] for g in agg.values())
print(f"step = launches {a}..{a + step_len - 1} ({step_len} launches, {tot_us:.1f} us serialised under ncu)")
for k, g in agg.items():
    print(f"{k:32s} n={g['launches']:3d} {g['us']:9.1f} us ({100 * g['us'] / tot_us:5.1f} %)  read {g['read'] / 1e6:9.1f} MB  "
          f"write {g['write'] / 1e6:9.1f} MB  -> {(g['read'] + g['write']) / g['us'] / 1e3:7.1f} GB/s")
conv = [g for k, g in agg.items() if k.startswith("conv_igemm_kernel")]
out = dict(source=sys.argv[1], step_launches=step_len,
           conv_stack_dram_bytes=sum(g["read"] + g["write"] for g in conv),
           conv_stack_launches=sum(g["launches"] for g in conv),
           conv_stack_share_of_step=sum(g["us"] for g in conv) / tot_us,
           per_kernel={k: g for k, g in agg.items()})
print(json.dumps({k: v for k, v in out.items() if k != "per_kernel"}))
if len(sys.argv) > 2:
    json.dump(out, open(sys.argv[2], "w"), indent=1)
