#!/bin/bash
# usage: tools/ncu_summary.sh file.ncu-rep > profiles/xxx_summary.txt   -- the few counters the notes cite
set -e
rep=$1
ncu -i "$rep" --page details --csv 2>/dev/null | python3 -c '
import csv,sys
keep=["Memory Throughput","DRAM Throughput","Duration","Compute (SM) Throughput","Executed Ipc Active","Issue Slots Busy","SM Busy",
"No Eligible","Eligible Warps Per Scheduler","Avg. Active Threads Per Warp","Registers Per Thread","Theoretical Occupancy","Achieved Occupancy",
"Dynamic Shared Memory Per Block","Static Shared Memory Per Block","Block Limit Registers","Block Limit Shared Mem","Executed Instructions","Kernel Name"]
r=csv.DictReader(sys.stdin); name=None
for row in r:
    if name is None: name=row.get("Kernel Name"); print("Kernel", name[:140])
    m=row.get("Metric Name","")
    if m in keep: print(m,row.get("Metric Unit",""),row.get("Metric Value",""))
'
ncu -i "$rep" --page raw --csv 2>/dev/null | python3 -c '
import csv,sys
rows=list(csv.reader(sys.stdin)); h=rows[0]; v=rows[-1]
d=dict(zip(h,v))
for k in sorted(d):
    if k.startswith("smsp__pcsamp_warps_issue_stalled") and not k.endswith("not_issued") : 
        pass
for k in ("sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active","sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active","sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active","sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active","sm__inst_executed_pipe_cbu.avg.pct_of_peak_sustained_active","sm__inst_executed_pipe_adu.avg.pct_of_peak_sustained_active","sm__inst_executed_pipe_fmaheavy.avg.pct_of_peak_sustained_active","sm__inst_executed_pipe_fmalite.avg.pct_of_peak_sustained_active","smsp__thread_inst_executed_per_inst_executed.ratio","dram__bytes.sum.per_second","sm__throughput.avg.pct_of_peak_sustained_elapsed"):
    if k in d: print("pipe/ratio", k, d[k])
want=["dram__bytes_read.sum","dram__bytes_write.sum","smsp__inst_executed.sum","sm__inst_executed_pipe_fma.sum","sm__inst_executed_pipe_alu.sum","sm__inst_executed_pipe_xu.sum","l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum","smsp__thread_inst_executed.sum"]
for k in want:
    if k in d: print(k, d[k])
st={k:float(d[k].replace(",","")) for k in d if k.startswith("smsp__pcsamp_warps_issue_stalled_") and "not_issued" not in k and d[k] not in ("","n/a")}
tot=sum(st.values()) or 1
print("stall samples (share):"," ".join("%s %.1f%%"%(k.replace("smsp__pcsamp_warps_issue_stalled_",""),100*x/tot) for k,x in sorted(st.items(),key=lambda t:-t[1])[:8]))
'
