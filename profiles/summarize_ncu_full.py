"""Summarise an `ncu --set full` report into profiles/<name>.{txt,json} (per-launch values of the metrics the
DESIGN/bench roofline cites).

    ncu -i gpurun_out/r1c_full.ncu-rep --page raw --csv > /tmp/raw.csv
    python profiles/summarize_ncu_full.py /tmp/raw.csv profiles/r1_ncu_full_summary "<command line that was profiled>"
"""
import csv
import json
import sys

METRICS = [
    "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
    "sm__throughput.avg.pct_of_peak_sustained_elapsed", "sm__warps_active.avg.pct_of_peak_sustained_active",
    "sm__pipe_tensor_subpipe_hmma_cycles_active.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_tensor_subpipe_hmma.avg.pct_of_peak_sustained_active",
    "sm__ops_path_tensor_op_hmma_src_bf16_dst_fp32_sparsity_off.sum",
    "sm__ops_path_tensor_op_hmma_src_bf16_dst_fp32_sparsity_off.sum.pct_of_peak_sustained_elapsed",
    "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum",
    "l1tex__data_pipe_lsu_wavefronts_mem_shared_op_ld.sum", "l1tex__data_pipe_lsu_wavefronts_mem_shared_op_st.sum",
    "l1tex__data_bank_conflicts_pipe_lsu_mem_shared_op_ld.sum", "l1tex__data_bank_conflicts_pipe_lsu_mem_shared_op_st.sum",
    "smsp__sass_inst_executed_op_shared_ld.sum", "smsp__sass_inst_executed_op_shared_st.sum",
    "memory_l1_wavefronts_shared", "memory_l1_wavefronts_shared_ideal", "derived__memory_l1_wavefronts_shared_excessive",
    "launch__registers_per_thread", "launch__grid_size",
    "launch__block_size", "launch__occupancy_limit_registers", "launch__occupancy_limit_shared_mem",
    "smsp__inst_executed.sum", "smsp__inst_executed_pipe_alu.sum", "smsp__inst_executed_pipe_fmaheavy.sum",
    "smsp__inst_executed_pipe_fmalite.sum", "smsp__inst_executed_pipe_fma.sum", "smsp__inst_executed_pipe_lsu.sum",
    "smsp__inst_executed_pipe_fp64.sum", "smsp__cycles_active.avg", "l1tex__t_sector_hit_rate.pct", "lts__t_sector_hit_rate.pct",
    "dram__throughput.avg.pct_of_peak_sustained_elapsed", "lts__throughput.avg.pct_of_peak_sustained_elapsed",
    "smsp__warp_issue_stalled_long_scoreboard_per_warp_active.pct", "smsp__issue_active.avg.pct_of_peak_sustained_active",
]


def num(s):
    try:
        return float(s.replace(",", ""))
    except ValueError:
        return None


def main():
    raw, out, cmd = sys.argv[1], sys.argv[2], (sys.argv[3] if len(sys.argv) > 3 else "")
    rows = list(csv.reader(open(raw)))
    # first row = metric names, second = units, then one row per profiled launch
    hdr = next(i for i, r in enumerate(rows) if "Kernel Name" in r)
    names, units = rows[hdr], rows[hdr + 1]
    kcol = names.index("Kernel Name")
    res = []
    for r in rows[hdr + 2:]:
        if len(r) != len(names):
            continue
        e = {"kernel": r[kcol]}
        for m in METRICS:
            if m in names:
                j = names.index(m)
                e[m] = {"value": num(r[j]), "unit": units[j]}
        res.append(e)
    json.dump(res, open(out + ".json", "w"), indent=1)
    with open(out + ".txt", "w") as f:
        f.write(f"# ncu --set full --clock-control none, {cmd}\n# per-launch values; raw report stays in gpurun_out/ (scratch)\n")
        for e in res:
            f.write("\n" + e["kernel"] + "\n")
            for m in METRICS:
                if m in e and e[m]["value"] is not None:
                    f.write(f"    {m:<100s} {e[m]['value']:>16.6f} {e[m]['unit']}\n")
    print(f"{len(res)} launches -> {out}.txt/.json")


if __name__ == "__main__":
    main()
