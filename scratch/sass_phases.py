"""Aggregate an `ncu --page source --csv` SASS export between barriers: samples / instructions per phase."""
import csv, sys
rows=list(csv.reader(open(sys.argv[1])))
hdr=[r for r in rows if r and r[0]=='Address'][0]
data=[r for r in rows if r and r[0].startswith('0x')]
si=hdr.index('# Samples'); ie=hdr.index('Instructions Executed'); src=hdr.index('Source')
tot=sum(int(r[si]) for r in data); toti=sum(int(r[ie]) for r in data)
print("total samples",tot,"total inst",toti, "n sass", len(data))
acc=0; acci=0; start=0
for k,r in enumerate(data):
    acc+=int(r[si]); acci+=int(r[ie])
    s=r[src].strip()
    if 'BAR' in s or s.startswith('EXIT') or k==len(data)-1:
        print("%5d..%5d  samples %6d (%4.1f%%) inst %9d (%4.1f%%)  %s"%(start,k,acc,100*acc/tot,acci,100*acci/toti,s[:40]))
        acc=0; acci=0; start=k+1
