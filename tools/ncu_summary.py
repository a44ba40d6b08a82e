#!/usr/bin/env python
"""Key counters of an `ncu --set full` report as a markdown table (one column per launch).

    python tools/ncu_summary.py gpurun_out/x.ncu-rep [> profiles/x.md]
"""
import csv
import subprocess
import sys

WANT = [
    ("gpu__time_duration.sum", "kernel time"),
    ("sm__cycles_elapsed.max", "SM cycles"),
    ("dram__bytes_read.sum", "DRAM read"),
    ("dram__bytes_write.sum", "DRAM write"),
    ("lts__t_sector_hit_rate.pct", "L2 hit rate"),
    ("l1tex__t_sector_hit_rate.pct", "L1 hit rate"),
    ("sm__ops_path_tensor_op_utchmma_src_fp16_dst_fp32_sparsity_off.avg.pct_of_peak_sustained_elapsed", "tensor pipe (fp16 UTCHMMA), % of peak"),
    ("sm__ops_path_tensor_op_utchmma_src_bf16_dst_fp32_sparsity_off.avg.pct_of_peak_sustained_elapsed", "tensor pipe (bf16 UTCHMMA), % of peak"),
    ("sm__mem_tensor_cycles_active.avg.pct_of_peak_sustained_active", "tensor memory path active, %"),
    ("sm__throughput.avg.pct_of_peak_sustained_elapsed", "SM throughput, %"),
    ("smsp__issue_active.avg.pct_of_peak_sustained_active", "issue slots active, %"),
    ("smsp__inst_executed.sum", "warp instructions"),
    ("smsp__warps_eligible.avg.per_cycle_active", "eligible warps / scheduler / clk"),
    ("l1tex__data_pipe_lsu_wavefronts_mem_shared.sum", "LSU shared-memory wavefronts"),
    ("l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "of which bank conflicts"),
    ("launch__registers_per_thread", "registers / thread"),
    ("launch__shared_mem_per_block_dynamic", "dynamic smem / CTA"),
    ("launch__grid_size", "grid"),
    ("launch__block_size", "block"),
    ("launch__cluster_size", "cluster"),
]


def main(path):
    out = subprocess.run(["ncu", "-i", path, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(out.splitlines()))
    hdr, units, data = rows[0], rows[1], rows[2:]
    idx = {h: i for i, h in enumerate(hdr)}
    names = [r[idx["Kernel Name"]].split("(")[0][-60:] for r in data]
    print("| metric | " + " | ".join(f"{i}: `{n}`" for i, n in enumerate(names)) + " |")
    print("|---|" + "---:|" * len(names))
    for key, label in WANT:
        if key not in idx:
            continue
        u = units[idx[key]]
        vals = []
        for r in data:
            v = r[idx[key]]
            try:
                f = float(v.replace(",", ""))
                v = f"{f:,.2f}" if abs(f) < 1e6 else f"{f:,.0f}"
            except ValueError:
                pass
            vals.append(v)
        print(f"| {label} ({u}) | " + " | ".join(vals) + " |")


if __name__ == "__main__":
    main(sys.argv[1])
