import sys, time, torch
sys.path.insert(0, '.')
from ubresnet_b200 import synthetic
from ubresnet_b200.deploy import wholeview as wv
from ubresnet_b200.models import UResNet
torch.manual_seed(1)
m = UResNet(3,1,16).cuda().eval()
pl = synthetic.make_wholeview().cuda()
def t(fn, n=10):
    for _ in range(3): fn()
    torch.cuda.synchronize(); e0=torch.cuda.Event(enable_timing=True); e1=torch.cuda.Event(enable_timing=True)
    t0=time.perf_counter(); e0.record()
    for _ in range(n): fn()
    host=(time.perf_counter()-t0)*1e3/n
    e1.record(); torch.cuda.synchronize(); return e0.elapsed_time(e1)/n, host
plan = wv.tile_plan(1008,3456,512)
crops = torch.cat([wv.gather(pl[p], plan, 512) for p in range(3)],0)
with torch.no_grad():
    print('event resident  gpu ms, host enqueue ms:', t(lambda: wv.run_event_classes(m, pl)))
    print('forward 42 crops:', t(lambda: m(crops)))
    print('forward 32 crops:', t(lambda: m(crops[:32])))
    lp = m(crops)
    from ubresnet_b200 import ops
    print('argmax:', t(lambda: ops.argmax_classes(lp)))
    print('gather x3 + cat:', t(lambda: torch.cat([wv.gather(pl[p], plan, 512) for p in range(3)],0)))
    cls = ops.argmax_classes(lp)
    print('stitch x3:', t(lambda: [wv.stitch_classes(cls[p*14:(p+1)*14], plan, 1008, 3456) for p in range(3)]))
