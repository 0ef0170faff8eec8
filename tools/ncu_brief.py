"""Brief of an .ncu-rep: python tools/ncu_brief.py rep [top]: hot SASS lines + key raw metrics."""
import csv, subprocess, sys, os
sys.path.insert(0, os.path.dirname(__file__))
import ncu_source_hot
rep = sys.argv[1]; top = int(sys.argv[2]) if len(sys.argv) > 2 else 25
subprocess.run(f"ncu -i {rep} --page source --csv > /tmp/_src.csv 2>/dev/null", shell=True)
subprocess.run(f"ncu -i {rep} --page raw --csv > /tmp/_raw.csv 2>/dev/null", shell=True)
ncu_source_hot.main("/tmp/_src.csv", top)
rows = list(csv.reader(open("/tmp/_raw.csv")))
hdr = rows[0]; vals = rows[2] if len(rows) > 2 else rows[1]
keys = ['sm__warps_active.avg.per_cycle_active', 'sm__inst_executed.avg.per_cycle_active', 'smsp__average_warps_issue_stalled',
        'smsp__issue_active.avg.pct', 'gpu__time_duration.sum', 'l1tex__data_bank_conflicts_pipe_lsu_mem_shared_op', 'launch__grid_size',
        'sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active', 'dram__bytes_read.sum ', 'dram__bytes_write.sum ',
        'l1tex__data_pipe_lsu_wavefronts_mem_shared.sum', 'smsp__inst_executed.sum']
for h, v in zip(hdr, vals):
    if any(k in h for k in keys) and not h.endswith('pct_of_peak_sustained_elapsed') and 'per_second' not in h:
        try:
            if float(v.replace(',', '')) == 0: continue
        except ValueError: pass
        print(h, '=', v)
