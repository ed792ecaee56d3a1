"""Summaries of ncu output for profiles/.
   python tools/ncu_summary.py launches gpurun_out/r2_launches.csv "<command>" > profiles/r2_launches_summary.txt
   ncu -i gpurun_out/r2_full.ncu-rep --page raw --csv > /tmp/raw.csv
   python tools/ncu_summary.py full /tmp/raw.csv "<command>" profiles/r2_ncu_full_summary      (writes .txt and .json)"""
import csv
import json
import sys

FULL_METRICS = [
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
    "launch__registers_per_thread", "launch__grid_size", "launch__block_size", "launch__occupancy_limit_registers",
    "launch__occupancy_limit_shared_mem", "smsp__inst_executed.sum", "smsp__cycles_active.avg",
    "l1tex__t_sector_hit_rate.pct", "lts__t_sector_hit_rate.pct", "lts__throughput.avg.pct_of_peak_sustained_elapsed",
    "smsp__issue_active.avg.pct_of_peak_sustained_active",
]


def rows_of(path):
    lines = [l for l in open(path, newline="") if not l.startswith("==")]
    return list(csv.reader(lines))


def launches(path, command):
    rows = rows_of(path)
    head = next(i for i, r in enumerate(rows) if "Kernel Name" in r)
    names = rows[head]
    kn, mn, mv, mu = names.index("Kernel Name"), names.index("Metric Name"), names.index("Metric Value"), names.index("Metric Unit")
    agg = {}
    n = 0
    for r in rows[head + 1:]:
        if len(r) <= mv or r[mn] != "gpu__time_duration.sum":
            continue
        v = float(r[mv].replace(",", ""))
        v *= {"ns": 1e-3, "us": 1.0, "ms": 1e3, "s": 1e6}.get(r[mu], 1.0)
        t, c = agg.get(r[kn], (0.0, 0))
        agg[r[kn]] = (t + v, c + 1)
        n += 1
    total = sum(t for t, _ in agg.values())
    print(f"# {path.replace('gpurun_out', 'profiles')}: {n} launches. {command}")
    print("# per-launch times under ncu are cold-cache and serialised: compare SHARES, not absolutes")
    for k, (t, c) in sorted(agg.items(), key=lambda kv: -kv[1][0]):
        print(f"{t:9.1f}us {100 * t / total:5.1f}% n={c:4d} avg={t / c:7.2f}us  {k[:150]}")
    print(f"total {total:.1f} us")


def full(path, command, out):
    rows = rows_of(path)
    head = next(i for i, r in enumerate(rows) if "Kernel Name" in r)
    names, units = rows[head], rows[head + 1]
    kn = names.index("Kernel Name")
    res = []
    with open(out + ".txt", "w") as f:
        f.write(f"# ncu --set full --clock-control none, {command}\n# per-launch values; raw report stays in gpurun_out/ (scratch)\n")
        for r in rows[head + 2:]:
            if len(r) <= kn:
                continue
            rec = {"kernel": r[kn]}
            f.write(f"\n{r[kn]}\n")
            for m in FULL_METRICS:
                if m not in names:
                    continue
                j = names.index(m)
                try:
                    v = float(r[j].replace(",", ""))
                except ValueError:
                    v = None
                rec[m] = {"value": v, "unit": units[j]}
                f.write(f"    {m:100s} {r[j]:>16s} {units[j]}\n")
            res.append(rec)
    json.dump(res, open(out + ".json", "w"), indent=1)


if __name__ == "__main__":
    if sys.argv[1] == "launches":
        launches(sys.argv[2], sys.argv[3])
    else:
        full(sys.argv[2], sys.argv[3], sys.argv[4])
