#!/bin/bash
# is the shipped tiling table limited by the noise of 10-launch timings?  rebuild the batch-64 entries with 40 launches per candidate, A/B
mkdir -p gpurun_out
TUNE_CHAINS="feed,c3" AQUA_TUNE_ITERS=40 timeout 900 python tools/make_tune_table.py gpurun_out/tune40.txt 64:640 > gpurun_out/tune40_r2am.log 2>&1; echo "tune exit=$?"; tail -1 gpurun_out/tune40_r2am.log
cat aquaculture_b200/tuned/b200.txt > /tmp/t40.txt; grep -v "^#" gpurun_out/tune40.txt > /tmp/t40_new.txt
python - <<'PY'
old = dict(l.rsplit(" ", 1) for l in open("/tmp/t40.txt") if l.strip() and not l.startswith("#"))
new = dict(l.rsplit(" ", 1) for l in open("/tmp/t40_new.txt") if l.strip())
diff = [k for k in new if k in old and old[k].strip() != new[k].strip()]
print(len(new), "entries,", len(diff), "choices differ")
merged = dict(old); merged.update(new)
open("/tmp/t40_merged.txt", "w").write("".join(f"{k} {v.strip()}\n" for k, v in merged.items()))
PY
AQUA_TUNE_LOG=0 timeout 900 python tools/ab_env.py - TABLE=/tmp/t40_merged.txt > gpurun_out/ab_tune40_r2am.txt 2> gpurun_out/ab_tune40_r2am.err; cat gpurun_out/ab_tune40_r2am.txt | cut -c1-170; tail -2 gpurun_out/ab_tune40_r2am.err
