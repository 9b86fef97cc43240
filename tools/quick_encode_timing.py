import torch, time, sys
import os; sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from audiocraft_b200.quantization import ResidualVectorQuantizer
torch.manual_seed(0)
m = ResidualVectorQuantizer(dimension=128, n_q=4, bins=2048, kmeans_init=False).cuda().eval()
m.vq.check_codes=False
xs=[torch.randn(64,128,1500,device='cuda') for _ in range(4)]
for algo in ['tc']:
    m.vq.encode_algo=algo
    for i in range(5): m.encode(xs[i%4])
    torch.cuda.synchronize()
    ts=[]
    for i in range(30):
        a=torch.cuda.Event(enable_timing=True); b=torch.cuda.Event(enable_timing=True)
        a.record(); c=m.encode(xs[i%4]); b.record(); torch.cuda.synchronize()
        ts.append(a.elapsed_time(b))
    ts.sort(); print(algo,'min %.3f med %.3f max %.3f ms'%(ts[0],ts[len(ts)//2],ts[-1]))
    # back-to-back
    a=torch.cuda.Event(enable_timing=True); b=torch.cuda.Event(enable_timing=True)
    a.record()
    for i in range(30): c=m.encode(xs[i%4])
    b.record(); torch.cuda.synchronize(); print('b2b per call %.3f ms'%(a.elapsed_time(b)/30))
    t0=time.perf_counter()
    for i in range(30): c=m.encode(xs[i%4])
    t1=time.perf_counter(); torch.cuda.synchronize(); print('host per call %.3f ms'%((t1-t0)/30*1e3))
    # encode timed with a decode (49 MB of streaming stores) between consecutive encodes
    ts=[]
    for i in range(30):
        a=torch.cuda.Event(enable_timing=True); b=torch.cuda.Event(enable_timing=True)
        a.record(); c=m.encode(xs[i%4]); b.record(); q=m.decode(c); torch.cuda.synchronize()
        ts.append(a.elapsed_time(b))
    ts.sort(); print(algo,'with decode between: min %.3f med %.3f max %.3f ms'%(ts[0],ts[len(ts)//2],ts[-1]))
    ts=[]
    for i in range(30):
        a=torch.cuda.Event(enable_timing=True); b=torch.cuda.Event(enable_timing=True)
        a.record(); c=m.encode(xs[i%4]); b.record(); q=m.decode(c)
        ts.append((a,b))
    torch.cuda.synchronize()
    ts=sorted(a.elapsed_time(b) for a,b in ts); print(algo,'async loop with decode: min %.3f med %.3f max %.3f ms'%(ts[0],ts[len(ts)//2],ts[-1]))
