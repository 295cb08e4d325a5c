"""Barrier-to-barrier segments of a kernel from an ncu --set full --import-source on capture.

  python profiles/segments.py gpurun_out/purchase_r1m.ncu-rep [min share of samples] [items per launch] > profiles/r01_purchase_segments.txt

The SASS page is cut at every BAR.SYNC; per segment: share of the warp instructions, instructions per item (CTA = one
purchasing manager / one seller), share of the stall samples with the leading stall reasons, and the hottest
instructions.  Samples are per warp, so a segment in which one warp works and the others wait at the next barrier shows
up twice: in its own samples and in the `barrier` samples of the segment that follows.
"""
import csv
import subprocess
import sys

rep=sys.argv[1]
out=subprocess.run(["ncu","-i",rep,"--page","source","--csv","--print-source","sass"],capture_output=True,text=True).stdout
rows=list(csv.reader(out.splitlines()))
hdr=rows[1]; iS=hdr.index("# Samples"); iE=hdr.index("Instructions Executed"); iSrc=hdr.index("Source")
st=[(i,h) for i,h in enumerate(hdr) if h.startswith('stall_') and 'Not Issued' not in h]
data=rows[2:]
tot_s=sum(int(r[iS] or 0) for r in data); tot_e=sum(int(r[iE] or 0) for r in data)
print("inst",tot_e,"samples",tot_s,"sass",len(data))
seg=[];cur=[]
for k,r in enumerate(data):
    cur.append((k,r))
    if "BAR.SYNC" in r[iSrc]:
        seg.append(cur);cur=[]
seg.append(cur)
thr=float(sys.argv[2]) if len(sys.argv)>2 else 0.015
items=int(sys.argv[3]) if len(sys.argv)>3 else 131274
for sg in seg:
    ss=sum(int(r[iS] or 0) for _,r in sg); ee=sum(int(r[iE] or 0) for _,r in sg)
    if ss<thr*tot_s: continue
    agg={h:sum(int(r[i] or 0) for _,r in sg) for i,h in st}
    top=sorted(agg.items(), key=lambda x:-x[1])[:6]
    print("seg sass[%d..%d] inst %.1f%% (%d per item) samples %.1f%%: "%(sg[0][0],sg[-1][0],100*ee/tot_e,ee/items,100*ss/tot_s)+", ".join("%s %.1f%%"%(h[6:],100*v/tot_s) for h,v in top))
    hot=sorted(sg,key=lambda kr:-int(kr[1][iS] or 0))[:4]
    for k,r in hot:
        print("      [%5d] %5.2f%%  %s"%(k,100*int(r[iS] or 0)/tot_s,r[iSrc].strip()[:90]))
agg={h:sum(int(r[i] or 0) for r in data) for i,h in st}
print({h[6:]:round(100*v/tot_s,1) for h,v in agg.items() if v>0.01*tot_s})
