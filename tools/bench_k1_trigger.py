import sys,os; sys.path.insert(0,".")
import torch, ssb200
H=W=513
def t(fn,n=30):
    for i in range(5): fn(i)
    torch.cuda.synchronize(); e0,e1=torch.cuda.Event(enable_timing=True),torch.cuda.Event(enable_timing=True); e0.record()
    for i in range(n): fn(i)
    e1.record(); torch.cuda.synchronize(); return e0.elapsed_time(e1)/n*1e3
g=torch.Generator(device="cuda"); g.manual_seed(0)
geom=ssb200.FrameGeometry(1920,1080,H,W)
for b in (64,256):
    cam=[torch.randint(0,256,(b,1080,1920,3),dtype=torch.uint8,device="cuda",generator=g) for _ in range(2)]
    rgb=torch.empty((b,H,W,3),dtype=torch.uint8,device="cuda")
    q=torch.randint(0,256,(b,H,W,21),dtype=torch.uint8,device="cuda",generator=g)
    lab=torch.empty((b,H,W),dtype=torch.uint8,device="cuda"); st=torch.empty((b,21,4),dtype=torch.int64,device="cuda")
    k1=min(t(lambda i: ssb200.preprocess_frames(cam[i%2],geom,out=rgb)) for _ in range(3))
    def both(i):
        ssb200.preprocess_frames(cam[i%2],geom,out=rgb); ssb200.argmax_stats(q,288,513,labels=lab,stats=st)
    c3=min(t(both,20) for _ in range(3))
    k2=min(t(lambda i: ssb200.argmax_stats(q,288,513,labels=lab,stats=st)) for _ in range(3))
    print("B=%d K1 %.1f us  K2 %.1f us  K1+K2 %.1f us"%(b,k1,k2,c3),flush=True)
    del cam,q
