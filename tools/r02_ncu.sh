#!/bin/bash
# Round-2 ncu evidence on one B200 (each capture only after the same command ran clean without ncu).
P="python tools/prof_case.py"
FULL="ncu --set full --clock-control none --import-source on -k regex:spmv_flag_kernel"
AB="ncu --clock-control none -k regex:spmv_flag_kernel --csv --metrics gpu__time_duration.sum,lts__t_sector_hit_rate.pct,dram__bytes_read.sum,dram__bytes_write.sum,lts__t_sectors_srcunit_tex_op_read_evict_last_lookup_hit.sum,lts__t_sectors_srcunit_tex_op_read_evict_last_lookup_miss.sum,lts__t_sectors_srcunit_tex_op_read_evict_normal_lookup_hit.sum,lts__t_sectors_srcunit_tex_op_read_evict_normal_lookup_miss.sum,l1tex__m_l1tex2xbar_req_cycles_active.avg.pct_of_peak_sustained_elapsed,dram__cycles_active.avg.pct_of_peak_sustained_elapsed"
o=gpurun_out
# full captures: pure band (1 slab), C3 as shipped (3 slabs), banded(11)+5 in 1 slab and in the 3 slabs setup picks
$P band16 kernel=3 > $o/pc1.log 2>&1 && $FULL -s 2 -c 1 -o $o/r02_band16_k3 $P band16 kernel=3 > $o/nc1.log 2>&1
$P c3 kernel=3 > $o/pc2.log 2>&1 && $FULL -s 6 -c 3 -o $o/r02_c3_k3 $P c3 kernel=3 > $o/nc2.log 2>&1
$P banded kernel=3 > $o/pc3.log 2>&1 && $FULL -s 2 -c 1 -o $o/r02_banded_k3 $P banded kernel=3 > $o/nc3.log 2>&1
$P banded3 kernel=3 > $o/pc8.log 2>&1 && $FULL -s 6 -c 3 -o $o/r02_banded3_k3 $P banded3 kernel=3 > $o/nc8.log 2>&1
# the L2 "vector cache" A/B north_star asks for: 3 slabs + access-policy window, 3 slabs without, 1 slab
$P c3 kernel=3 persist=1 > $o/pc4.log 2>&1 && $AB -s 6 -c 3 --log-file $o/r02_l2ab_3slabs_window.csv $P c3 kernel=3 persist=1 > $o/nc4.log 2>&1
$P c3 kernel=3 persist=0 > $o/pc5.log 2>&1 && $AB -s 6 -c 3 --log-file $o/r02_l2ab_3slabs_nopersist.csv $P c3 kernel=3 persist=0 > $o/nc5.log 2>&1
$P c3 kernel=3 slabs=1 persist=1 > $o/pc6.log 2>&1 && $AB -s 2 -c 1 --log-file $o/r02_l2ab_1slab.csv $P c3 kernel=3 slabs=1 persist=1 > $o/nc6.log 2>&1
# every launch of a short bench run
B="python bench.py --steps 5 --warmup 3 --no-cpu --no-others --no-parity"
$B > $o/pc7.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -c 800 --csv --log-file $o/r02_launches.csv $B > $o/nc7.log 2>&1
for f in pc1 pc2 pc3 pc8 pc4 pc5 pc6 pc7; do echo "$f: $(cut -c1-300 $o/$f.log)"; done
