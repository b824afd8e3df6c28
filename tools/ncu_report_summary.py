"""Text summary of an .ncu-rep (ncu --set full --import-source on): headline metrics, warp-stall mix and issued-instruction mix per captured kernel.
Usage: python tools/ncu_report_summary.py <report.ncu-rep> <out.txt> "<title>"   (used for profiles/ncu_*_r2_*.txt)"""
import csv, subprocess, sys, collections
rep, out, title = sys.argv[1], sys.argv[2], sys.argv[3]
raw = subprocess.run(["ncu","-i",rep,"--page","raw","--csv"],capture_output=True,text=True).stdout
rows=list(csv.reader(raw.splitlines())); hdr=rows[0]; units=rows[1]
want=["Kernel Name","launch__grid_size","launch__block_size","launch__registers_per_thread","gpu__time_duration.sum","dram__bytes_read.sum","dram__bytes_write.sum","gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed","sm__throughput.avg.pct_of_peak_sustained_elapsed","l1tex__throughput.avg.pct_of_peak_sustained_elapsed","lts__throughput.avg.pct_of_peak_sustained_elapsed","smsp__inst_executed.sum","sm__warps_active.avg.pct_of_peak_sustained_active","smsp__issue_active.avg.pct_of_peak_sustained_active","sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active","l1tex__t_sector_hit_rate.pct","lts__t_sector_hit_rate.pct","smsp__thread_inst_executed_per_inst_executed.ratio"]
src = subprocess.run(["ncu","-i",rep,"--page","source","--csv"],capture_output=True,text=True).stdout
blocks=[]; cur=None
for r in csv.reader(src.splitlines()):
    if r and r[0]=="Kernel Name": cur={"name":r[1],"hdr":None,"rows":[]}; blocks.append(cur); continue
    if cur is None: continue
    if cur["hdr"] is None: cur["hdr"]=r; continue
    cur["rows"].append(r)
with open(out,"w") as f:
    f.write(title+"\n")
    f.write("ncu --set full --clock-control none --import-source on (one launch per kernel; cold-cache replay: durations are not bench values)\n\n")
    for k,r in enumerate(rows[2:]):
        for w in want:
            if w in hdr:
                i=hdr.index(w); f.write(f"{w} = {r[i]} [{units[i]}]\n")
        if k < len(blocks):
            b=blocks[k]; h=b["hdr"]; ia=h.index("Source"); ie=h.index("Instructions Executed")
            st=collections.Counter(); byop=collections.Counter(); tot=0
            sc=[(i,x) for i,x in enumerate(h) if x.startswith("stall_") and "Not Issued" not in x]
            for rr in b["rows"]:
                try: e=int(rr[ie])
                except: continue
                t=rr[ia].split()
                if not t: continue
                op=(t[1] if t[0].startswith("@") and len(t)>1 else t[0]).split(".")[0]
                byop[op]+=e; tot+=e
                for i,x in sc:
                    try: st[x]+=int(rr[i])
                    except: pass
            stot=sum(st.values()) or 1
            f.write("warp stall samples: "+", ".join(f"{x[6:]} {c/stot*100:.0f}%" for x,c in st.most_common(6))+"\n")
            f.write("issued instruction mix: "+", ".join(f"{op} {c/tot*100:.1f}%" for op,c in byop.most_common(10))+"\n")
        f.write("\n")
