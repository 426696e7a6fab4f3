"""Keep the judged columns of an `ncu --page raw --csv` export: python tools/ncu_trim.py raw.csv out.csv"""
import csv
import re
import sys

rows = list(csv.reader(open(sys.argv[1])))
hdr = rows[0]
pat = (r'gpu__time_duration|dram__bytes_(read|write)\.sum$|gpu__dram_throughput|sm__pipe_tensor|sm__warps_active|'
       r'launch__registers|launch__shared_mem|launch__occupancy_limit|smsp__issue_active|sm__throughput|lts__t_bytes\.sum$|'
       r'lts__t_sector_hit_rate|l1tex__data_pipe_lsu_wavefronts_mem_shared\.sum$|smsp__inst_executed\.sum$|'
       r'sm__cycles_elapsed\.avg$|smsp__average_warps_issue_stalled.*_per_issue_active')
keep = [i for i, h in enumerate(hdr) if h in ('ID', 'Kernel Name', 'Block Size', 'Grid Size') or re.search(pat, h)]
with open(sys.argv[2], 'w', newline='') as f:
    w = csv.writer(f)
    for r in rows:
        w.writerow([r[i] for i in keep])
print(len(keep), 'columns,', len(rows) - 2, 'launches')
