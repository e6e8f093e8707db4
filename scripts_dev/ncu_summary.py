"""One line per captured launch of an .ncu-rep (ncu --set full): the metrics the design notes quote.
    python scripts_dev/ncu_summary.py gpurun_out/x.ncu-rep > profiles/r2_x.csv"""
import csv, io, re, subprocess, sys
WANT = [("gpu__time_duration.sum", "time_us", "time"), ("launch__grid_size", "grid", 1), ("launch__block_size", "block", 1),
        ("launch__registers_per_thread", "regs", 1), ("launch__shared_mem_per_block_dynamic", "dyn_smem_KB", 1),
        ("sm__throughput.avg.pct_of_peak_sustained_elapsed", "sm_pct", 1),
        ("sm__issue_active.avg.pct_of_peak_sustained_elapsed", "issue_active_pct", 1),
        ("sm__inst_executed_pipe_tensor_subpipe_hmma.avg.pct_of_peak_sustained_active", "hmma_pct", 1),
        ("sm__pipe_tensor_subpipe_hmma_cycles_active.avg.pct_of_peak_sustained_active", "hmma_cycles_pct", 1),
        ("sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active", "alu_pct", 1),
        ("sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active", "fma_pct", 1),
        ("sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active", "lsu_pct", 1),
        ("smsp__warps_active.avg.per_cycle_active", "warps_per_sched", 1),
        ("smsp__inst_executed.sum", "warp_insts", 1),
        ("dram__bytes_read.sum", "dram_rd_MB", None), ("dram__bytes_write.sum", "dram_wr_MB", None),
        ("dram__throughput.avg.pct_of_peak_sustained_elapsed", "dram_pct", 1),
        ("lts__t_sector_hit_rate.pct", "l2_hit_pct", 1),
        ("l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "smem_bank_conflicts", 1)]
UNIT = {"byte": 1e-6, "Kbyte": 1e-3, "Mbyte": 1.0, "Gbyte": 1e3}
TIME = {"ns": 1e-3, "us": 1.0, "ms": 1e3, "s": 1e6, "nsecond": 1e-3, "usecond": 1.0, "msecond": 1e3, "second": 1e6}
out = subprocess.run(["ncu", "-i", sys.argv[1], "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
hdr, units = rows[0], rows[1]
ix = {h: i for i, h in enumerate(hdr)}
w = csv.writer(sys.stdout)
w.writerow(["kernel"] + [n for _, n, _ in WANT if _ in ix or True])
for r in rows[2:]:
    name = re.sub(r"\(anonymous namespace\)::|<unnamed>::|void ", "", r[ix["Kernel Name"]]).split("(")[0]
    line = [name]
    for m, _, scale in WANT:
        if m not in ix:
            line.append("")
            continue
        v = float(r[ix[m]].replace(",", "")) if r[ix[m]] not in ("", "n/a") else float("nan")
        if scale == "time":
            v *= TIME.get(units[ix[m]], 1.0)
        elif scale is None:
            v *= UNIT.get(units[ix[m]], 1.0)
        else:
            v *= scale
        line.append(f"{v:.4g}")
    w.writerow(line)
