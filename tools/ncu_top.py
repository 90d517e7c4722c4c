import csv, sys, subprocess
rep=sys.argv[1]; top=int(sys.argv[2]) if len(sys.argv)>2 else 45
out=subprocess.run(['ncu','-i',rep,'--page','source','--csv','--print-source','sass'],capture_output=True,text=True).stdout
rows=list(csv.reader(out.splitlines()))
hdr=rows[1]
ia=hdr.index('Address'); isrc=hdr.index('Source'); iex=hdr.index('Instructions Executed'); ismp=hdr.index('# Samples')
ist=hdr.index('Warp Stall Sampling (All Samples)')
data=[]
for r in rows[2:]:
    try: data.append((int(r[ia],16), int(r[iex]), int(r[ismp]), r[isrc]))
    except: pass
data.sort()
tot=sum(d[2] for d in data); totex=sum(d[1] for d in data)
print('kernel', rows[0][1][:90]); print('total samples', tot, 'total instr', totex)
idx={a:i for i,(a,_,_,_) in enumerate(data)}
for a,ex,smp,s in sorted(data,key=lambda d:-d[2])[:top]:
    i=idx[a]; prev=data[i-1][3].strip()[:46] if i>0 else ''
    print(f"{smp:5d} {100*smp/tot:4.1f}% ex={ex:8d} {a&0xfffff:05x} {s.strip()[:58]:58s} | prev: {prev}")
