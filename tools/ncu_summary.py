import sys,csv,subprocess
rep=sys.argv[1]
out=subprocess.run(["ncu","-i",rep,"--page","raw","--csv"],capture_output=True,text=True).stdout
rows=list(csv.reader(out.splitlines()))
hdr,units,vals=rows[0],rows[1],rows[2]
d={h:(v,u) for h,v,u in zip(hdr,vals,units)}
keys=["Kernel Name","gpu__time_duration.sum","launch__grid_size","launch__block_size","launch__registers_per_thread","launch__shared_mem_per_block_dynamic",
"sm__throughput.avg.pct_of_peak_sustained_elapsed","gpu__compute_memory_throughput.avg.pct_of_peak_sustained_elapsed",
"dram__throughput.avg.pct_of_peak_sustained_elapsed","dram__bytes_read.sum","dram__bytes_write.sum",
"lts__throughput.avg.pct_of_peak_sustained_elapsed","l1tex__throughput.avg.pct_of_peak_sustained_elapsed",
"sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active","sm__inst_executed_pipe_tensor.sum","smsp__issue_active.avg.pct_of_peak_sustained_active",
"sm__warps_active.avg.pct_of_peak_sustained_active","lts__t_sector_hit_rate.pct","lts__t_bytes.sum","l1tex__data_pipe_lsu_wavefronts_mem_shared.sum"]
for k in keys:
    if k in d: print(f"{k}: {d[k][0]} {d[k][1]}")
print("--- stall reasons (warps per issue)")
st=[(h,float(v.replace(',',''))) for h,(v,u) in d.items() if h.startswith("smsp__average_warps_issue_stalled") and h.endswith("_per_issue_active.ratio") and v not in ("","n/a")]
for h,v in sorted(st,key=lambda x:-x[1])[:8]: print(f"{h.replace('smsp__average_warps_issue_stalled_','').replace('_per_issue_active.ratio','')}: {v:.2f}")
for h,(v,u) in d.items():
    if "tma" in h.lower() or "tmem" in h.lower(): print(h,v,u)
