import sys, torch
sys.path.insert(0,'/root/repo'); sys.path.insert(0,'/root/repo/tests')
from msgat_b200 import _lib, models, synthetic
from test_gpu_parity import _model_and_batch
model,batch=_model_and_batch('pemsd8',3)
model=model.cuda()
X=batch.X.cuda()
tpc=model.tpcs[0]
x0=X[:,0].contiguous()
m1,m2=tpc.tgacns
rel=lambda a,b: float((a.double()-b.double()).norm()/b.double().norm())
with torch.no_grad():
    h1=m1(x0,model.adj)
    o32=m2(h1,model.adj)
    m2.precision=_lib.PREC_TF32
    otc=m2(h1,model.adj)
    m2.precision=0
print('layer2 out rel',rel(otc,o32))
print('mask flips', int(((otc>0)!=(o32>0)).sum()), 'of', otc.numel(), 'frac zero', float((o32==0).float().mean()))
d=(otc-o32).abs()
print('max abs diff',float(d.max()),'out absmax',float(o32.abs().max()), 'mean abs', float(o32.abs().mean()))
for name,sl in (('cacn',slice(0,24)),('tacn',slice(24,48)),('gacn',slice(48,72))):
    print(name, rel(otc[:,sl],o32[:,sl]))
# per node error
pn=(otc-o32).double().pow(2).sum((0,1,3)).sqrt()/o32.double().pow(2).sum((0,1,3)).sqrt()
print('per-node rel err: max',float(pn.max()),'argmax',int(pn.argmax()),'median',float(pn.median()))
# gradient check of layer-2 alone with random upstream
torch.manual_seed(1)
R=torch.randn_like(o32)
def grads(prec):
    m2.precision=prec
    for p in m2.parameters(): p.grad=None
    hh=h1.clone().requires_grad_(True)
    o=m2(hh,model.adj); (o*R).sum().backward()
    return hh.grad.clone(), {n:p.grad.clone() for n,p in m2.named_parameters()}
g32,p32=grads(0); gtc,ptc=grads(_lib.PREC_TF32)
print('dx rel',rel(gtc,g32))
rows=sorted(((rel(ptc[n],p32[n]),n) for n in p32),reverse=True)
for r in rows[:8]: print(r)
