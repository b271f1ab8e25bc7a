"""Compact summary of an ncu --set full report (raw page): python tools/ncu_summary.py file.ncu-rep [more.ncu-rep ...]
Prints, per captured launch, the metrics the roofline in bench.py / DESIGN.md quotes (duration, DRAM bytes,
tensor-pipe activity, L2/L1 throughput, launch geometry)."""
import csv
import subprocess
import sys

WANT = [
    "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "dram__throughput.avg.pct_of_peak_sustained_elapsed",
    "lts__t_bytes.sum", "lts__throughput.avg.pct_of_peak_sustained_elapsed", "l1tex__throughput.avg.pct_of_peak_sustained_elapsed",
    # tensor path: the *_utchmma_* counters are the ones tcgen05.mma (SASS UTCHMMA) increments -- flops executed and their share of
    # the 8192 flop/clk/SM peak; the legacy *_hmma_* / pipe_tensor_cycles_active counters stay ~0 for UTCHMMA kernels
    "sm__ops_path_tensor_op_utchmma_src_fp16_dst_fp32_sparsity_off.sum",
    "sm__ops_path_tensor_op_utchmma_src_fp16_dst_fp32_sparsity_off.avg.pct_of_peak_sustained_elapsed",
    "sm__inst_executed_pipe_tensor.sum", "sm__mem_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed",
    "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_elapsed", "sm__issue_active.avg.pct_of_peak_sustained_elapsed",
    "l1tex__data_pipe_tc_wavefronts_mem_shared.sum.pct_of_peak_sustained_elapsed",
    "sm__cycles_active.avg", "sm__cycles_elapsed.avg", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
    "launch__grid_size", "launch__block_size", "launch__registers_per_thread", "launch__shared_mem_per_block_dynamic",
    "smsp__cycles_active.avg", "sm__warps_active.avg.pct_of_peak_sustained_active",
]
for rep in sys.argv[1:]:
    out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(out.splitlines()))
    H, U = rows[0], rows[1]
    for V in rows[2:]:
        d = dict(zip(H, zip(U, V)))
        print(f"== {rep}: {d.get('Kernel Name', ('', '?'))[1][:100]}")
        for k in WANT:
            for h in H:
                if h == k or h.endswith("." + k):
                    u, v = d[h]
                    if v != "":
                        print(f"   {k:85s} {v:>16s} {u}")
                    break
