#!/usr/bin/env python
"""Key metrics of every kernel in an .ncu-rep (read with `ncu -i ... --page raw --csv`) as the text
summaries committed under profiles/.   usage: ncu_summary.py report.ncu-rep > profiles/name.txt"""
import csv, subprocess, sys

KEYS = [
    "gpu__time_duration.sum", "launch__grid_size", "launch__block_size", "launch__registers_per_thread",
    "launch__shared_mem_per_block_dynamic", "sm__cycles_elapsed.max", "smsp__inst_executed.sum",
    "smsp__thread_inst_executed_per_inst_executed.ratio", "smsp__issue_active.avg.pct_of_peak_sustained_active",
    "sm__warps_active.avg.pct_of_peak_sustained_active", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
    "dram__bytes_read.sum", "dram__bytes_write.sum", "lts__t_sectors.sum", "lts__t_sector_hit_rate.pct",
    "lts__throughput.avg.pct_of_peak_sustained_elapsed", "l1tex__t_sector_hit_rate.pct",
    "l1tex__throughput.avg.pct_of_peak_sustained_active", "l1tex__data_pipe_lsu_wavefronts.avg.pct_of_peak_sustained_elapsed",
    "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum", "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum",
    "l1tex__data_pipe_tc_wavefronts_mem_shared.sum", "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active",
    "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed",
    "sm__ops_path_tensor_op_utchmma_src_fp16_dst_fp32_sparsity_off.avg.pct_of_peak_sustained_elapsed",
    "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active", "sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active",
]
STALLS = "smsp__pcsamp_warps_issue_stalled_"


def main():
    out = subprocess.run(["ncu", "-i", sys.argv[1], "--page", "raw", "--csv"], stdout=subprocess.PIPE, text=True).stdout
    rows = list(csv.reader(out.splitlines()))
    H, U = rows[0], rows[1]
    for r in rows[2:]:
        print("==", r[H.index("Kernel Name")][:90])
        for k in KEYS:
            if k in H:
                i = H.index(k)
                print(f"  {k} [{U[i]}] = {r[i]}")
        stalls = [(float(r[i] or 0), h[len(STALLS):]) for i, h in enumerate(H)
                  if h.startswith(STALLS) and "not_issued" not in h and r[i] not in ("", "n/a")]
        tot = sum(s for s, _ in stalls) or 1.0
        top = sorted(stalls, reverse=True)[:6]
        print("  warp-stall samples: " + ", ".join(f"{n} {s / tot * 100:.0f}%" for s, n in top))


if __name__ == "__main__":
    main()
