#!/usr/bin/env python
"""Summarise an `ncu --page source --csv` dump: stall mix, instruction classes, SASS listing."""
import csv, sys
rows=list(csv.reader(open(sys.argv[1])))
h=rows[1]; data=rows[2:]
ix={k:i for i,k in enumerate(h)}
stalls=[k for k in h if k.startswith('stall_') and 'Not Issued' not in k]
tot={k:0 for k in stalls}; samples=0; inst=0
for r in data:
    for k in stalls:
        try: tot[k]+=int(r[ix[k]])
        except: pass
    samples+=int(r[ix['# Samples']] or 0); inst+=int(r[ix['Instructions Executed']] or 0)
print('total samples',samples,'inst',inst)
for k,v in sorted(tot.items(), key=lambda x:-x[1])[:6]: print(k,v, '%.1f%%'%(100*v/max(samples,1)))
if len(sys.argv)>2 and sys.argv[2]=='list':
    lo=int(sys.argv[3]) if len(sys.argv)>3 else 0
    for r in data:
        n=int(r[ix['Instructions Executed']] or 0)
        if n>=lo: print(r[ix['# Samples']].rjust(6), str(n).rjust(9), r[ix['Source']][:100])
