#!/usr/bin/env python
"""Print selected metrics from an `ncu --page raw --csv` dump: ncu_keys.py raw.csv [substr ...]"""
import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
hdr, units = rows[0], rows[1]
DEFAULT = ['Kernel Name', 'gpu__time_duration.sum', 'dram__bytes_read.sum', 'dram__bytes_write.sum',
           'lts__t_sector_hit_rate.pct', 'sm__warps_active.avg.pct_of_peak_sustained_active',
           'launch__registers_per_thread', 'launch__grid_size', 'launch__block_size',
           'gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed',
           'l1tex__throughput.avg.pct_of_peak_sustained_elapsed', 'lts__throughput.avg.pct_of_peak_sustained_elapsed',
           'sm__throughput.avg.pct_of_peak_sustained_elapsed', 'smsp__issue_active.avg.pct_of_peak_sustained_active',
           'smsp__inst_executed.sum', 'sm__cycles_elapsed.max',
           'l1tex__t_sectors_pipe_lsu_mem_global_op_ld.sum', 'l1tex__t_requests_pipe_lsu_mem_global_op_ld.sum',
           'l1tex__data_pipe_lsu_wavefronts.sum', 'l1tex__data_pipe_lsu_wavefronts_mem_shared.sum',
           'l1tex__t_output_wavefronts_pipe_lsu_mem_global_op_ld.sum', 'lts__t_sectors_srcunit_tex_op_read.sum',
           'lts__t_sectors_op_read.sum', 'lts__t_sectors_op_write.sum']
subs = sys.argv[2:]
for r in rows[2:]:
    print('----')
    for i, k in enumerate(hdr):
        if (k in DEFAULT and not subs) or any(s in k for s in subs):
            print(k, '=', r[i][:100], units[i])
