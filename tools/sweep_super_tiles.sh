for cfg in "16 32" "32 32" "32 64" "41 32" "41 16" "82 16" "16 16" "21 128" "11 128"; do
  set -- $cfg
  PI0_SUPER_M=$1 PI0_SUPER_N=$2 ncu --metrics dram__bytes_read.sum,gpu__time_duration.sum,lts__t_sector_hit_rate.pct --clock-control none --kernel-name-base mangled -k regex:gemm_tc_pair_kernelILi256ELi1E -s 4 -c 3 --csv python tools/profile_step.py 64 1 2>/dev/null | python -c "
import sys,csv
rows=[r for r in csv.reader(sys.stdin) if len(r)>5 and r[0].isdigit()]
import collections
d=collections.defaultdict(list)
for r in rows: d[r[-3]].append(float(r[-1].replace(',','')))
print('super', '$1', '$2', {k:(round(sum(v)/len(v),3)) for k,v in d.items()})"
done
