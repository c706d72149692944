import re,sys,csv,subprocess,collections
cubin,kern,srccsv=sys.argv[1],sys.argv[2],sys.argv[3]
out=subprocess.run(['nvdisasm','--print-line-info',cubin],capture_output=True,text=True).stdout
# split into functions
lines=out.splitlines()
cur=None; curline=None; amap={}
for l in lines:
    m=re.match(r'\s*\.text\.(\S+):',l)
    if m: cur=m.group(1); continue
    if cur is None or kern not in cur: continue
    m=re.search(r'//## File "([^"]+)", line (\d+)(.*)',l)
    if m:
        curline=(m.group(1).split('/')[-1],int(m.group(2)), 'inl' in m.group(3)); continue
    m=re.match(r'\s*/\*([0-9a-f]{4,})\*/\s+(.*?);',l)
    if m: amap[int(m.group(1),16)]=(curline,m.group(2))
rows=list(csv.reader(open(srccsv)))
hdr=rows[1]; ia=hdr.index('Address'); isamp=hdr.index('# Samples'); iex=hdr.index('Instructions Executed'); ith=hdr.index('Thread Instructions Executed')
base=None; agg=collections.defaultdict(lambda:[0,0,0])
tot=[0,0,0]
for r in rows[2:]:
    try: a=int(r[ia],16)
    except: continue
    if base is None: base=a
    off=a-base
    ln=amap.get(off,(None,''))[0]
    key=(ln[0],ln[1]) if ln else ('?',0)
    s=int(r[isamp] or 0); e=int(r[iex] or 0); t=int(r[ith] or 0)
    for k in (key,):
        agg[k][0]+=s; agg[k][1]+=e; agg[k][2]+=t
    tot[0]+=s; tot[1]+=e; tot[2]+=t
print("total samples %d inst %d thread-inst %d avg lanes %.1f"%(tot[0],tot[1],tot[2],tot[2]/max(tot[1],1)))
for k,v in sorted(agg.items(),key=lambda x:-x[1][1])[:int(sys.argv[4]) if len(sys.argv)>4 else 45]:
    print("%-18s:%4d  samples %5.1f%%  inst %5.1f%%  lanes %4.1f"%(k[0],k[1],100*v[0]/tot[0],100*v[1]/tot[1],v[2]/max(v[1],1)))
