import sys,os; sys.path.insert(0,".")
import numpy as np, torch, ssb200
g=torch.Generator(device="cuda"); g.manual_seed(11)
def tables(mask,lab,rh,rw,rows,path):
    os.environ["SSB_BLOB_PATH"]=path
    t=ssb200.blob_regions(mask,lab,rh,rw,21,rows); torch.cuda.synchronize()
    return {k:v.cpu().numpy() for k,v in t.items()}
def same(a,c,rows):
    ok=(a["count"]==c["count"]).all()
    for f in range(len(a["count"])):
        n=int(a["count"][f])
        if n>rows: continue
        oa,oc=np.argsort(a["first"][f,:n]),np.argsort(c["first"][f,:n])
        for k in ("first","parent","flags","hist"): ok=ok and (a[k][f,:n][oa]==c[k][f,:n][oc]).all()
        holes=(a["flags"][f,:n][oa]&3)==0
        ok=ok and (a["ring"][f,:n][oa][holes]==c["ring"][f,:n][oc][holes]).all()
    return bool(ok)
for (B,H,W,rh,rw,rows) in ((300,200,257,200,257,512),(1,513,513,513,513,4096),(150,96,640,96,640,300),(77,513,513,288,500,256),(9,700,720,700,720,2000)):
    lab=torch.randint(0,21,(B,H,W),dtype=torch.uint8,device="cuda",generator=g)
    cells=torch.randint(0,4,(B,H//8+1,W//8+1),device="cuda",generator=g)
    blob=torch.tensor([0,7,15,3],device="cuda",dtype=torch.uint8)[cells].repeat_interleave(8,1).repeat_interleave(8,2)[:,:H,:W].contiguous()
    lab[::2]=blob[::2]
    mask=ssb200.foreground_mask(lab,rh,rw)
    a=tables(mask,lab,rh,rw,rows,"cluster"); c=tables(mask,lab,rh,rw,rows,"global")
    print(B,H,W,rh,rw,rows,"regions/frame %.0f"%a["count"].mean(),"same:",same(a,c,rows),flush=True)
