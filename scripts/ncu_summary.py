"""Selected raw metrics of an ncu report as a small CSV (one column per captured launch):
    python scripts/ncu_summary.py gpurun_out/r1_fused_c4.ncu-rep > profiles/r1_ncu_full_c4_fused.csv"""
import csv
import subprocess
import sys

KEEP = ["dram__bytes_read.sum", "dram__bytes_write.sum", "dram__sectors_read.sum", "dram__sectors_write.sum",
        "gpc__cycles_elapsed.avg.per_second", "gpc__cycles_elapsed.max", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
        "gpu__time_duration.sum", "l1tex__t_sectors_pipe_lsu_mem_local_op_ld.sum", "launch__registers_per_thread",
        "lts__t_sector_hit_rate.pct", "lts__t_sectors_srcunit_tex_op_read.sum",
        "lts__t_sectors_srcunit_tex_op_read_lookup_miss.sum", "lts__t_sectors_srcunit_tex_op_write.sum",
        "lts__throughput.avg.pct_of_peak_sustained_elapsed", "sm__cycles_active.avg",
        "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active",
        "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "smsp__average_warp_latency_per_inst_issued.ratio", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
        "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum", "smsp__inst_executed.sum",
        # bytes arriving at the SMs from L2 (the operand-delivery bound of the BF16-correction kernels)
        "l1tex__m_xbar2l1tex_read_bytes.sum", "l1tex__m_xbar2l1tex_read_bytes.sum.per_second",
        "l1tex__m_xbar2l1tex_read_bytes.sum.pct_of_peak_sustained_elapsed", "derived__lts__lts2xbar_bytes.sum.per_second",
        "lts__lts2xbar_cycles_active.avg.pct_of_peak_sustained_elapsed"]
out = subprocess.run(["ncu", "-i", sys.argv[1], "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(l for l in out.splitlines() if l.startswith('"')))
hdr, units, data = rows[0], rows[1], rows[2:]
w = csv.writer(sys.stdout)
w.writerow(["metric", "unit"] + [r[hdr.index("Kernel Name")] for r in data])
for i, h in enumerate(hdr):
    if h in KEEP:
        w.writerow([h, units[i]] + [r[i] for r in data])
