import csv,sys,subprocess
rep,kern,skip=sys.argv[1],sys.argv[2],sys.argv[3]
out=subprocess.run(['ncu','-i',rep,'--page','source','--csv','--print-source','cuda,sass','-k','regex:'+kern,'--launch-skip',skip,'--launch-count','1'],capture_output=True,text=True).stdout
rows=list(csv.reader(out.splitlines()))
secs=[];cur=None;fname=''
for r in rows:
    if len(r)==2 and r[0] in('File Path','File Name'): fname=r[1].split('/')[-1]
    if len(r)>5 and r[0]=='Line No': cur={'h':r,'rows':[],'f':fname}; secs.append(cur)
    elif cur is not None and len(r)==len(cur['h']): cur['rows'].append(r)
allr=[]
for s in secs:
    h=s['h']; ie=h.index('Instructions Executed'); ws=h.index('Warp Stall Sampling (All Samples)')
    for r in s['rows']:
        if not r[0]: continue
        try: n=int(r[ie]); st=int(r[ws] or 0)
        except: continue
        if n or st: allr.append((s['f'],r[0],r[1].strip()[:100],n,st))
tot=sum(a[3] for a in allr); tots=sum(a[4] for a in allr)
print('total warp inst',tot,'stall samples',tots)
for a in sorted(allr,key=lambda a:-(a[3]/tot+a[4]/tots))[:int(sys.argv[4]) if len(sys.argv)>4 else 30]:
    print(f"{a[0][:18]:18s}:{a[1]:>4} inst {a[3]/tot*100:5.1f}% stall {a[4]/tots*100:5.1f}%  {a[2]}")
