import sys, time; sys.path.insert(0,'.'); sys.path.insert(0,'tests')
import numpy as np, torch
import orbslam2_b200 as ob
import bench_configs as BC
import os; cfg = BC.CFG[int(os.environ.get("CFG","2"))]; n=int(os.environ.get("NPAIRS","64"))
left,right = BC.synth_inputs(cfg, 0, n)
p = BC.params_for(cfg)
exL, exR = ob.ORBextractor(**p), ob.ORBextractor(**p)
cap = exL.max_keypoints(cfg["w"], cfg["h"]); print("cap", cap)
hL, hR = torch.from_numpy(left).pin_memory(), torch.from_numpy(right).pin_memory()
def outs():
    return (torch.empty((n,cap,7),dtype=torch.float32).pin_memory(), torch.empty((n,cap,32),dtype=torch.uint8).pin_memory())
oL,oR = outs(),outs()
def T(fn, k=10):
    fn(); torch.cuda.synchronize(); t0=time.perf_counter()
    for _ in range(k): fn()
    torch.cuda.synchronize(); return (time.perf_counter()-t0)/k*1e3
def left_only():
    exL.submit_batch(hL,oL); exL.wait_batch()
def both():
    exL.submit_batch(hL,oL); exR.submit_batch(hR,oR); exL.wait_batch(); exR.wait_batch()
def full():
    both(); ob.compute_stereo_matches(exL,exR,386.1,0.0,718.8)
print("left only ms", T(left_only)); print("both ms", T(both)); print("both+stereo ms", T(full))
for ch in (8,16,32):
    exL.set_chunk(ch); exR.set_chunk(ch)
    print("chunk",ch,"both ms", T(both), "full", T(full))
dL, dR = hL.cuda(), hR.cuda()
dOL = (torch.empty((n,cap,7),dtype=torch.float32,device='cuda'), torch.empty((n,cap,32),dtype=torch.uint8,device='cuda'))
dOR = (torch.empty_like(dOL[0]), torch.empty_like(dOL[1]))
exL.set_chunk(0); exR.set_chunk(0)
def dev():
    exL.submit_batch(dL,dOL); exR.submit_batch(dR,dOR); exL.wait_batch(); exR.wait_batch()
def devfull():
    dev(); ob.compute_stereo_matches(exL,exR,386.1,0.0,718.8)
print("device both ms", T(dev), "device full", T(devfull))

for ch in (0, n, max(n // 4, 1)):
    exL.set_chunk(ch); exR.set_chunk(ch)
    print("device chunk", ch, "both ms", T(dev), "full", T(devfull))
