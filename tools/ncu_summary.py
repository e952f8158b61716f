#!/usr/bin/env python
"""Prints the metrics DESIGN.md / profiles/ quote from an .ncu-rep (run where ncu is installed; no GPU needed)."""
import csv, subprocess, sys
WANT = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
        "lts__t_sector_hit_rate.pct", "l1tex__t_sector_hit_rate.pct", "l1tex__throughput.avg.pct_of_peak_sustained_active",
        "sm__throughput.avg.pct_of_peak_sustained_elapsed", "smsp__issue_active.avg.pct_of_peak_sustained_active",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "smsp__thread_inst_executed_per_inst_executed.ratio",
        "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_fmaheavy.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_fmalite.avg.pct_of_peak_sustained_active", "sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_tex.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_uniform.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_cbu.avg.pct_of_peak_sustained_active",
        "launch__registers_per_thread", "launch__grid_size", "launch__occupancy_limit_registers", "sm__cycles_elapsed.avg",
        "smsp__inst_executed.sum", "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum"]
rows = list(csv.reader(subprocess.run(["ncu", "-i", sys.argv[1], "--page", "raw", "--csv"], capture_output=True, text=True).stdout.splitlines()))
hdr, units, data = rows[0], rows[1], rows[2:]
print("kernels:", [r[hdr.index("Kernel Name")][:50] for r in data])
extra = [h for h in hdr if h not in WANT and ("issue_stalled" in h and "per_issue_active" in h or "pipe" in h and ".avg.pct_of_peak_sustained_active" in h)]
extra += [h for h in hdr if h not in WANT and h in ("l1tex__data_pipe_lsu_wavefronts.sum", "l1tex__data_pipe_tex_wavefronts.sum", "l1tex__t_requests_pipe_lsu_mem_global_op_ld.sum",
          "l1tex__t_sectors_pipe_lsu_mem_global_op_ld.sum", "l1tex__t_requests_pipe_tex.sum", "l1tex__lsu_writeback_active.avg.pct_of_peak_sustained_active",
          "l1tex__tex_writeback_active.avg.pct_of_peak_sustained_active", "l1tex__data_bank_reads.avg.pct_of_peak_sustained_elapsed", "lts__t_sectors_srcunit_tex.sum",
          "l1tex__f_wavefronts.sum", "l1tex__m_xbar2l1tex_read_sectors.sum", "sm__inst_executed_pipe_fmaheavy.avg.pct_of_peak_sustained_active")]
for w in WANT + extra:
    if w in hdr:
        i = hdr.index(w)
        print("%-90s %-14s %s" % (w, units[i], "  ".join(r[i] for r in data)))
