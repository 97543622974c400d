import sys
rows=[l.split() for l in open(sys.argv[1]) if l.startswith('launch')]
ev=[(int(r[1]),int(r[3]),float(r[5])) for r in rows]
lo=int(sys.argv[2]) if len(sys.argv)>2 else 28
cur={}; out=[]; names={3:'cn',5:'fin',10:'repack',11:'retire'}
tot={}
for i,k,ms in ev:
    tot[k]=tot.get(k,0)+ms
    if k==31: cur['vn']=ms; out.append(cur); cur={}
    elif k in names: cur[names[k]]=cur.get(names[k],0)+ms
for j,c in enumerate(out):
    if j>=lo: print(j, {k:round(v,3) for k,v in c.items()}, round(sum(c.values()),3))
print(cur); print({k:round(v,2) for k,v in tot.items()}, round(sum(tot.values()),2))
