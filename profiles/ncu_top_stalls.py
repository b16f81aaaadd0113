"""Top stall-sampled SASS instructions of a kernel:  ncu -i rep --page source --csv > src.csv; python profiles/ncu_top_stalls.py src.csv [N]"""
import csv, sys
rows=list(csv.reader(open(sys.argv[1])))
# find header row
hi=[i for i,r in enumerate(rows) if r and r[0]=="Address"]
start=hi[0]; end=hi[1]-1 if len(hi)>1 else len(rows)
hdr=rows[start]
ia=hdr.index("Source"); isamp=hdr.index("# Samples"); iex=hdr.index("Instructions Executed")
body=rows[start+1:end]
tot=sum(int(r[isamp]) for r in body if len(r)>isamp and r[isamp].isdigit())
print("total samples",tot, "instructions", len(body))
order=sorted(range(len(body)), key=lambda i:-int(body[i][isamp]) if body[i][isamp].isdigit() else 0)
for i in order[:int(sys.argv[2]) if len(sys.argv)>2 else 25]:
    r=body[i]
    print(f"--- idx {i} samples {r[isamp]} ({100*int(r[isamp])/tot:.1f}%) exec {r[iex]}")
    for j in range(max(0,i-3), min(len(body), i+3)):
        print(("  >> " if j==i else "     ")+body[j][ia].strip()[:110], body[j][isamp])
