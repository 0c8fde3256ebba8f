import csv,sys
rows=list(csv.reader(open(sys.argv[1])))
cur_file=None; data=[]
for r in rows:
    if r and r[0]=='File Path': cur_file=r[1].split('/')[-1]; continue
    if len(r)>9 and r[0].isdigit():
        try: s=int(r[6]); i=int(r[7]); t=int(r[8])
        except: continue
        data.append((s,i,t,cur_file,r[0],r[1].strip()[:105]))
tot_s=sum(d[0] for d in data); tot_i=sum(d[1] for d in data); tot_t=sum(d[2] for d in data)
print("samples",tot_s,"warp-instr",tot_i,"avg lanes %.1f"%(tot_t/tot_i))
N=int(sys.argv[2]) if len(sys.argv)>2 else 40
for s,i,t,f,ln,src in sorted(data,key=lambda x:-x[0])[:N]:
    print("%5.1f%% smp %5.1f%% inst lanes %4.1f | %s:%s %s"%(100*s/tot_s,100*i/tot_i,t/max(i,1),f,ln,src))
