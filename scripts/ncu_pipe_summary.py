"""Shared-memory data-pipe summary of one kernel from an `ncu --page raw --csv` dump (K7t analysis: LSU + tensor
wavefronts against elapsed cycles).  usage: python scripts/ncu_pipe_summary.py raw.csv [slabs]"""
import csv
import sys

rows = list(csv.reader(open(sys.argv[1])))
hdr, units, r = rows[0], rows[1], rows[-1]
slabs = float(sys.argv[2]) if len(sys.argv) > 2 else None


def val(k):
    return float(r[hdr.index(k)].replace(",", "")) if k in hdr and r[hdr.index(k)] not in ("", "n/a") else float("nan")


keys = [
    "gpu__time_duration.sum",
    "sm__cycles_elapsed.avg",
    "sm__pipe_tensor_cycles_active_realtime.avg.pct_of_peak_sustained_elapsed",
    "TPC.TriageCompute.sm__pipe_tensor_cycles_active_realtime.avg.pct_of_peak_sustained_elapsed",
    "l1tex__data_pipe_lsu_wavefronts.avg.pct_of_peak_sustained_elapsed",
    "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum",
    "l1tex__data_pipe_lsu_wavefronts_mem_shared_op_ld.sum",
    "l1tex__data_pipe_lsu_wavefronts_mem_shared_op_st.sum",
    "l1tex__data_pipe_tc_wavefronts_mem_shared.sum",
    "SM_A.TriageCompute.l1tex__data_pipe_lsu_wavefronts.avg",
    "SM_A.TriageCompute.l1tex__data_pipe_lsu_wavefronts_mem_lgds.avg",
    "l1tex__data_bank_conflicts_pipe_lsu_mem_shared_op_ld.sum",
    "l1tex__data_bank_conflicts_pipe_lsu_mem_shared_op_st.sum",
    "l1tex__m_xbar2l1tex_read_bytes.sum",
    "l1tex__m_xbar2l1tex_read_bytes_mem_global_op_tma_ld.sum",
    "smsp__issue_active.avg.pct_of_peak_sustained_active",
    "sm__warps_active.avg.pct_of_peak_sustained_active",
    "launch__registers_per_thread",
    "dram__bytes_read.sum",
]
print("kernel:", r[hdr.index("Kernel Name")][:60])
for k in keys:
    if k in hdr:
        v = val(k)
        extra = ""
        if slabs and k.endswith(".sum") and "wavefronts" in k or (slabs and "conflicts" in k):
            extra = f"   = {v / slabs:8.1f} per k-slab MMA"
        if slabs and k.endswith("lsu_wavefronts.avg") or (slabs and k.endswith("lgds.avg")):
            extra = f"   = {v * 148 / slabs:8.1f} per k-slab MMA (x148 SMs)"
        print(f"{k:95s} {units[hdr.index(k)]:8s} {v:16.2f}{extra}")
