import csv,sys,subprocess
rep=sys.argv[1]
out=subprocess.run(["ncu","-i",rep,"--page","raw","--csv"],capture_output=True,text=True).stdout
rows=list(csv.reader(out.splitlines()))
hdr=rows[0]; units=rows[1]
pats=sys.argv[2:] or ['gpu__time_duration.sum','dram__bytes_read.sum','dram__bytes_write.sum','sm__throughput.avg.pct','gpu__dram_throughput.avg.pct','lts__t_bytes.sum','lts__throughput','sm__warps_active.avg.pct','launch__registers','pipe_tensor','sm__inst_executed_pipe_alu.avg.pct','sm__inst_executed_pipe_fma.avg.pct','sm__inst_executed_pipe_uniform','smsp__issue_active.avg.pct','sm__cycles_elapsed.avg.per_second','issue_stalled.*per_issue_active','l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum','sm__cycles_active.avg']
import re
for i,h in enumerate(hdr):
    if any(re.search(p,h) for p in pats):
        print(h, units[i], [r[i] for r in rows[2:]])
