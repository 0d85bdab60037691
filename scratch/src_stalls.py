"""Per-CUDA-line stall attribution from `ncu --page source --csv --print-source cuda,sass`."""
import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
topn = int(sys.argv[2]) if len(sys.argv) > 2 else 40
cur_file=None; hdr=None; out=[]
def I(x):
    try: return int(x)
    except: return 0
for r in rows:
    if not r: continue
    if r[0]=='File Path': cur_file=r[1].split('/')[-1]; continue
    if r[0]=='Function Name': continue
    if r[0]=='Line No': hdr=r; continue
    if r[0] and r[0].isdigit():
        out.append((cur_file,int(r[0]),r[1],r))
iS=hdr.index('# Samples'); iI=hdr.index('Instructions Executed')
stall_cols=[i for i,h in enumerate(hdr) if h.startswith('stall_') and 'Not Issued' not in h]
tot=sum(I(x[3][iS]) for x in out); print('total',tot)
for f,l,s,r in sorted(out,key=lambda x:-I(x[3][iS]))[:topn]:
    st = sorted(((I(r[i]),hdr[i][6:]) for i in stall_cols), reverse=True)[:3]
    print(f[:14].ljust(14), str(l).rjust(4), r[iS].rjust(6), r[iI].rjust(9), s.strip()[:80].ljust(80), st)
