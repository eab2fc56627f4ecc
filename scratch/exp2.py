import os, sys, torch
sys.path.insert(0,'/root/repo'); sys.path.insert(0,'/root/repo/tests')
from msgat_b200 import _lib, models, synthetic
from test_gpu_parity import _model_and_batch
name=sys.argv[1] if len(sys.argv)>1 else 'pemsd4'
model,batch=_model_and_batch(name,2)
model=model.cuda(); X=batch.X.cuda()
rel=lambda a,b: float((a.double()-b.double()).norm()/b.double().norm())
for rep in range(2):
  for m,tpc in enumerate(model.tpcs):
    x0=X[:,m].contiguous()
    with torch.no_grad():
        h1=tpc.tgacns[0](x0,model.adj)
        os.environ['MSGAT_TC_DISABLE']='1'
        o32=tpc.tgacns[1](h1,model.adj)
        os.environ['MSGAT_TC_DISABLE']='0'
        otc=tpc.tgacns[1](h1,model.adj)
        torch.cuda.synchronize()
    d=(otc-o32).abs()
    idx=d.flatten().argmax()
    bb,cc,nn,tt=[int(v) for v in torch.unravel_index(idx,d.shape)]
    pern=d.amax((0,1,3))
    bad=(pern>1e-4).nonzero().flatten().tolist()
    print(rep,m,'rel',rel(otc,o32),'maxabs',float(d.max()),'at b,c,n,t',bb,cc,nn,tt,'bad nodes',bad[:20],len(bad))
