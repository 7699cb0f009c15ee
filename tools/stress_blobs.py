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
# label maps with long uniform stretches (the whole-word path of the label counts), cell sizes 8..100, label maps that
# start at every byte phase, ROIs narrower than the frame
for i, (B,H,W,rh,rw,rows,cell) in enumerate(((64,513,513,288,513,1024,16),(64,513,513,288,513,1024,40),(40,513,513,513,513,2048,100),
                                              (33,300,515,300,515,1024,32),(50,257,390,200,333,700,24),(20,120,1030,120,1030,900,64),
                                              (16,513,513,288,513,1024,600),(31,97,64,97,64,300,8),(25,64,31,64,31,100,5))):
    off=i%4
    buf=torch.zeros(B*H*W+8,dtype=torch.uint8,device="cuda")
    lab=buf[off:off+B*H*W].view(B,H,W)
    cells=torch.randint(0,21,(B,H//cell+2,W//cell+2),device="cuda",generator=g)
    cells[torch.rand(cells.shape,device="cuda",generator=g)<0.5]=15
    cells[torch.rand(cells.shape,device="cuda",generator=g)<0.3]=0
    lab.copy_(cells.repeat_interleave(cell,1).repeat_interleave(cell,2)[:,3:H+3,5:W+5].to(torch.uint8))
    lab[:, ::37, ::29]=7                       # single odd pixels inside the uniform stretches
    mask=ssb200.foreground_mask(lab,rh,rw)
    a=tables(mask,lab,rh,rw,rows,"cluster"); c=tables(mask,lab,rh,rw,rows,"global")
    print("cells",cell,B,H,W,rh,rw,rows,"phase",off,"regions/frame %.0f"%a["count"].mean(),"same:",same(a,c,rows),flush=True)
