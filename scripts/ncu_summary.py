"""Summarise an ncu report (one kernel launch) into a small JSON: python scripts/ncu_summary.py in.ncu-rep out.json [note]"""
import csv, io, json, subprocess, sys

KEEP = ('gpu__time_duration.sum', 'sm__cycles_elapsed.max', 'dram__bytes_read.sum', 'dram__bytes_write.sum', 'lts__t_sector_hit_rate.pct',
        'lts__t_bytes.sum', 'sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active', 'sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active',
        'sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active', 'sm__pipe_alu_cycles_active.avg.pct_of_peak_sustained_active',
        'smsp__issue_active.avg.pct_of_peak_sustained_active', 'l1tex__data_pipe_lsu_wavefronts_mem_shared.sum',
        'l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum', 'launch__registers_per_thread', 'launch__grid_size', 'launch__block_size',
        'launch__shared_mem_per_block_dynamic', 'sm__warps_active.avg.pct_of_peak_sustained_active', 'smsp__inst_executed.sum',
        'smsp__sass_inst_executed_op_tmem_ldt.sum', 'smsp__sass_inst_executed_op_tmem_stt.sum', 'dram__throughput.avg.pct_of_peak_sustained_elapsed',
        'gpu__compute_memory_throughput.avg.pct_of_peak_sustained_elapsed', 'sm__throughput.avg.pct_of_peak_sustained_elapsed')

rep, out = sys.argv[1], sys.argv[2]
note = sys.argv[3] if len(sys.argv) > 3 else ""
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr, units = rows[0], rows[1]
res = []
for vals in rows[2:]:
    d = {"kernel": vals[hdr.index("Kernel Name")] if "Kernel Name" in hdr else "?"}
    for i, h in enumerate(hdr):
        if h in KEEP or ("pcsamp_warps_issue_stalled" in h and "not_issued" not in h):
            d[h] = {"value": vals[i], "unit": units[i]}
    res.append(d)
json.dump({"report": rep, "note": note, "launches": res}, open(out, "w"), indent=1)
print(out, len(res), "launch(es)")
