import csv, collections, sys
rows=list(csv.reader(open(sys.argv[1])))
want=sys.argv[2] if len(sys.argv)>2 else 'kernels_fused'
sections=[]; cur=None
i=0
while i < len(rows):
    r=rows[i]
    if r and r[0]=="File Path":
        cur={'file':r[1],'func':rows[i+1][1],'hdr':rows[i+2],'rows':[]}; sections.append(cur); i+=3; continue
    if cur is not None and r: cur['rows'].append(r)
    i+=1
for sec in sections:
    hdr=sec['hdr']; n=len(hdr)
    if want not in sec['file']: continue
    idx={}
    for j,h in enumerate(hdr): idx.setdefault(h,j-n)      # from the end
    src_rows=[r for r in sec['rows'] if r[0]!='']
    S=idx['# Samples']; I=idx['Instructions Executed']
    def num(r,c):
        try: return int(r[c] or 0)
        except: return 0
    tot=sum(num(r,S) for r in src_rows); toti=sum(num(r,I) for r in src_rows)
    print("==", sec['file'].split('/')[-1], sec['func'][18:70], "samples", tot, "instr", toti)
    stall_cols=[h for h in hdr if h.startswith('stall_') and 'Not Issued' not in h]
    top=sorted(src_rows, key=lambda r:-num(r,S))[:int(sys.argv[3]) if len(sys.argv)>3 else 22]
    for r in top:
        st=sorted(((num(r,idx[c]),c) for c in stall_cols), reverse=True)[:3]
        print("%5s %6d smp %5.1f%% inst %9s | %s | %s" % (r[0], num(r,S), 100*num(r,S)/max(tot,1), r[I], ','.join(r[1:len(r)+idx['Address']]).strip()[:80], ' '.join('%s=%d'%(c[6:],v) for v,c in st)))
