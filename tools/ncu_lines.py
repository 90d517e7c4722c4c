"""Per-source-line instruction counts and stall samples from an .ncu-rep (needs -lineinfo + --import-source on)."""
import csv, io, subprocess, sys

def main(path, top=40, sort='inst'):
    out = subprocess.run(['ncu', '-i', path, '--page', 'source', '--csv', '--print-source', 'cuda,sass'], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    cur_file = None; hdr = None; recs = []
    for r in rows:
        if len(r) == 2 and r[0] == 'File Path': cur_file = r[1].split('/')[-1]; continue
        if r and r[0] == 'Line No': hdr = r; continue
        if hdr and len(r) == len(hdr) and r[0] != '':
            rec = dict(zip(hdr, r)); 
            try:
                inst = int(rec['Instructions Executed']); samp = int(rec['# Samples'])
            except ValueError:
                continue
            stalls = {h[6:]: int(rec[h]) for h in hdr if h.startswith('stall_') and '(' not in h and rec[h].isdigit()}
            top2 = sorted(stalls.items(), key=lambda kv: -kv[1])[:3]
            src = r[1]
            recs.append((cur_file, rec['Line No'], inst, samp, top2, src.strip()[:110]))
    tot_i = sum(r[2] for r in recs); tot_s = sum(r[3] for r in recs)
    print(f"total inst {tot_i}  samples {tot_s}")
    key = (lambda r: -r[2]) if sort == 'inst' else (lambda r: -r[3])
    for f, ln, inst, samp, st, src in sorted(recs, key=key)[:top]:
        print(f"{f}:{ln:>4} inst={inst/tot_i*100:5.1f}% samp={samp/tot_s*100:5.1f}% {','.join(f'{k}:{v}' for k,v in st if v):40s} | {src}")

if __name__ == '__main__':
    main(sys.argv[1], int(sys.argv[2]) if len(sys.argv) > 2 else 40, sys.argv[3] if len(sys.argv) > 3 else 'inst')
