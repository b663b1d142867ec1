import csv,collections,re,sys
rows=[r for r in csv.reader(open(sys.argv[1])) if len(r)>10 and r[0].isdigit()]
t=collections.Counter(); rd=collections.Counter(); wr=collections.Counter(); n=collections.Counter(); occ=collections.defaultdict(list); hit=collections.defaultdict(list)
for r in rows:
    name=re.sub(r'\(.*','',r[4]).split('::')[-1]; met=r[-3]; unit=r[-2]; v=float(r[-1].replace(',',''))
    if 'time' in met: t[name]+= v/1e6 if unit=='ns' else v/1e3 if unit=='us' else v; n[name]+=1
    elif 'warps_active' in met: occ[name].append(v)
    elif 'hit_rate' in met: hit[name].append(v)
    else:
        f={'byte':1,'Kbyte':1e3,'Mbyte':1e6,'Gbyte':1e9}[unit]
        (rd if 'read' in met else wr)[name]+=v*f
tot=sum(t.values()); print('total ms',tot, dict(n))
for k in t: print('%-22s %7.2f ms  read %6.2f GB write %6.2f GB -> %5.0f GB/s  occ %4.1f%% l2hit %4.1f%%'%(k,t[k],rd[k]/1e9,wr[k]/1e9,(rd[k]+wr[k])/t[k]/1e6,sum(occ[k])/len(occ[k]),sum(hit[k])/len(hit[k])))
