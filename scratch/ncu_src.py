import csv, sys
rows=list(csv.reader(open(sys.argv[1])))
hdr=rows[1]; data=[r for r in rows[2:] if len(r)==len(hdr)]
iS=hdr.index("# Samples"); iSrc=hdr.index("Source"); iEx=hdr.index("Instructions Executed")
stalls=[i for i,h in enumerate(hdr) if h.startswith('stall_') and 'Not Issued' not in h]
tot=sum(int(r[iS]) for r in data if r[iS].isdigit())
print('total samples',tot, 'instr rows',len(data))
N=int(sys.argv[2]) if len(sys.argv)>2 else 40
top=sorted([(int(r[iS]),k) for k,r in enumerate(data) if r[iS].isdigit()],reverse=True)[:N]
for s,k in sorted(top,key=lambda x:x[1]):
    r=data[k]
    st={hdr[i][6:]:int(r[i]) for i in stalls if r[i].isdigit() and int(r[i])>0}
    st=sorted(st.items(),key=lambda x:-x[1])[:3]
    print(k,s,f"{100*s/tot:.1f}%",r[iEx],r[iSrc].strip()[:60],st)
