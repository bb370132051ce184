import csv,collections,re,sys
for fn in sys.argv[1:]:
    lines=[l for l in open(fn) if not l.startswith('==')]
    r=csv.DictReader(lines)
    agg=collections.OrderedDict()
    for x in r:
        k=re.sub(r'\(.*','',x['Kernel Name'])
        v=float(x['Metric Value'].replace(',',''))
        a=agg.setdefault(k,{'n':0,'t':0.0,'i':0.0})
        if x['Metric Name']=='gpu__time_duration.sum':
            u=x['Metric Unit']
            if u=='ns': v/=1e6
            elif u=='us': v/=1e3
            elif u=='s': v*=1e3
            a['t']+=v; a['n']+=1
        elif x['Metric Name']=='smsp__inst_executed.sum': a['i']+=v
    tot=sum(a['t'] for a in agg.values())
    print(fn, "total %.2f ms"%tot)
    for k,a in sorted(agg.items(), key=lambda t:-t[1]['t'])[:7]: print("  %-40s n=%3d  %8.3f ms  %5.1f%%  inst %.3e"%(k[:40],a['n'],a['t'],100*a['t']/tot,a['i']))
