import csv, sys, subprocess
rep=sys.argv[1]; tiles=float(sys.argv[2])
out=subprocess.run(['ncu','-i',rep,'--page','source','--csv','--print-source','sass'],capture_output=True,text=True).stdout
rows=list(csv.reader(out.splitlines()))
hdr=rows[1]
ia=hdr.index('Address'); isrc=hdr.index('Source'); iex=hdr.index('Instructions Executed'); ismp=hdr.index('# Samples') if '# Samples' in hdr else None
data=[]
for r in rows[2:]:
    try: data.append((int(r[ia],16), int(r[iex]), int(r[ismp]) if ismp is not None else 0, r[isrc]))
    except: pass
data.sort()
tot=sum(d[1] for d in data)
regs=[]; cur=None
for a,ex,smp,s in data:
    if cur and abs(cur[1]-ex)<=0.02*max(ex,1): cur[2]+=1; cur[3]+=smp; cur[4]=a; cur[6]+=ex
    else:
        if cur: regs.append(cur)
        cur=[a,ex,1,smp,a,s,ex]
regs.append(cur)
print('instr per tile total', tot/tiles)
for a,ex,n,smp,a2,s,sumex in regs:
    if sumex/tot>0.01:
        print(f"{a&0xfffff:05x}-{a2&0xfffff:05x} x{ex/tiles:7.2f}/tile n={n:4d} instr/tile={sumex/tiles:7.1f} ({100*sumex/tot:4.1f}%) smp={smp:5d} {s.strip()[:50]}")
