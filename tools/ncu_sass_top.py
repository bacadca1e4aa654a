import csv, subprocess, sys, io
rep=sys.argv[1]; key=sys.argv[2] if len(sys.argv)>2 else "stall_long_sb"
src = subprocess.run(["ncu","-i",rep,"--page","source","--csv","--print-source","sass"],capture_output=True,text=True).stdout
hdr=None; rows=[]
for r in csv.reader(io.StringIO(src)):
    if r and r[0]=="Address": hdr=r; continue
    if hdr and len(r)==len(hdr) and r[0].startswith("0x"):
        rows.append(r)
ki=hdr.index(key); si=hdr.index("# Samples") if "# Samples" in hdr else 2
print([h for h in hdr][:12])
tot=sum(int(r[ki] or 0) for r in rows)
idx={r[0]:i for i,r in enumerate(rows)}
for r in sorted(rows,key=lambda r:-int(r[ki] or 0))[:14]:
    i=idx[r[0]]
    print(f"{100*int(r[ki])/max(tot,1):5.1f}% {r[0]} {r[1].strip()[:70]}   | prev: {rows[i-1][1].strip()[:50]}")
