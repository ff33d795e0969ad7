import csv,sys,subprocess
rep=sys.argv[1]
raw=subprocess.run(['ncu','-i',rep,'--page','raw','--csv'],capture_output=True,text=True).stdout
rows=list(csv.reader(raw.splitlines())); hdr,units,vals=rows[0],rows[1],rows[2]
want=["gpu__time_duration.sum","dram__bytes_read.sum","dram__bytes_write.sum","sm__throughput.avg.pct_of_peak_sustained_elapsed","l1tex__throughput.avg.pct_of_peak_sustained_elapsed","lts__throughput.avg.pct_of_peak_sustained_elapsed","dram__throughput.avg.pct_of_peak_sustained_elapsed","sm__warps_active.avg.pct_of_peak_sustained_active","launch__registers_per_thread","launch__occupancy_limit_shared_mem","launch__occupancy_limit_registers","l1tex__data_pipe_lsu_wavefronts.sum","l1tex__data_pipe_lsu_wavefronts_mem_shared.sum","l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum","l1tex__data_bank_conflicts_pipe_lsu_mem_shared_op_ld.sum","l1tex__data_bank_conflicts_pipe_lsu_mem_shared_op_st.sum","smsp__issue_active.avg.pct_of_peak_sustained_active","l1tex__data_pipe_lsu_wavefronts.avg.pct_of_peak_sustained_elapsed","sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active","smsp__inst_executed.sum","sm__cycles_elapsed.avg","launch__shared_mem_per_block_dynamic"]
for i,h in enumerate(hdr):
    if h in want: print("%-70s %-10s %s"%(h,units[i],vals[i]))
    if h.startswith("smsp__average_warps_issue_stalled") and h.endswith("per_issue_active.ratio"):
        try:
            if float(vals[i])>0.15: print("   stall %-50s %s"%(h[34:-23],vals[i]))
        except: pass
src=subprocess.run(['ncu','-i',rep,'--page','source','--csv'],capture_output=True,text=True).stdout
rows=list(csv.reader(src.splitlines()))
hi=[i for i,r in enumerate(rows) if r and r[0]=="Address"][0]
hdr=rows[hi]; ia=hdr.index("Source"); isamp=hdr.index("# Samples"); iex=hdr.index("Instructions Executed")
data=rows[hi+1:]
tot=sum(int(r[isamp]) for r in data); totex=sum(int(r[iex]) for r in data)
bars=[k for k,r in enumerate(data) if "BAR.SYNC" in r[ia]]+[len(data)]
prev=0
for b in bars:
    s=sum(int(r[isamp]) for r in data[prev:b]); e=sum(int(r[iex]) for r in data[prev:b])
    print("region [%4d,%4d) samples %5.1f%% inst %5.1f%% (%.0fM)"%(prev,b,100*s/tot,100*e/totex,e/1e6)); prev=b
thr=float(sys.argv[2]) if len(sys.argv)>2 else 0.8
for k,r in enumerate(data):
    s=int(r[isamp])
    if s>tot*thr/100: print("%4d %5.2f%% %9s  %s"%(k,100*s/tot,r[iex],r[ia].strip()[:90]))
