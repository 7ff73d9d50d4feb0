"""Diagnostic: per-kernel GPU timeline of one vector-API decode step (torch.profiler / CUPTI)."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from torch.profiler import profile, ProfilerActivity
from sca_ldpc_b200 import QaryDecoder
from sca_ldpc_b200.workloads import KyberWorkload

wl = KyberWorkload(seed=0, variant="kyber512", sum_weight=6, check_blocks=1, accuracy=0.98, nbits=4, nbits_sum=5)
T = 10000
b = wl.generate(T, seed=1000)
dev = torch.device("cuda", 0)
dec = QaryDecoder(wl.H, 10, device=0, **wl.decoder_kwargs)
d_in = [torch.from_numpy(b.priors).to(dev), torch.from_numpy(b.priors_sum).to(dev)]
ob = wl.generate_observed(T, seed=1000)
d_obs = [torch.from_numpy(ob.obs.view(np.int16)).to(dev), torch.from_numpy(ob.obs_sum.view(np.int16)).to(dev)]
d_tab = [torch.from_numpy(wl.table).to(dev), torch.from_numpy(wl.table_sum).to(dev)]

def vec():
    return dec.min_sum_batch(d_in[0], d_in[1], want_posterior=False)[0]
def obs():
    return dec.min_sum_observed(d_tab[0], d_obs[0], d_tab[1], d_obs[1])[0]

for name, fn in (("vec", vec), ("obs", obs), ("vec", vec)):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(10):
        fn()
    e1.record()
    torch.cuda.synchronize()
    print(name, "ms/step", e0.elapsed_time(e1) / 10)

with profile(activities=[ProfilerActivity.CUDA, ProfilerActivity.CPU]) as prof:
    vec()
    torch.cuda.synchronize()
evs = sorted([e for e in prof.events() if e.device_type == torch.autograd.DeviceType.CUDA], key=lambda e: e.time_range.start)
t0 = evs[0].time_range.start
prev_end = t0
for e in evs:
    print(f"{(e.time_range.start - t0):9.1f} us  gap {(e.time_range.start - prev_end):7.1f}  dur {e.time_range.elapsed_us():8.1f}  {e.name[:60]}")
    prev_end = e.time_range.end

# ---- the bench's step_dev (decode + success counting) ----------------------------------------------------
secret = torch.from_numpy(b.secret).to(dev)
nssum = torch.from_numpy((-b.ssum).astype(np.int8)).to(dev)
bv = wl.bvars
def step_dev():
    hard = vec()
    ok = (hard[:, :bv] == secret).all(1) & (hard[:, bv:] == nssum).all(1)
    return torch.stack((ok.sum(), torch.tensor(T, device=dev)))
def step_dev2():
    hard = vec()
    ok = (hard[:, :bv] == secret).all(1) & (hard[:, bv:] == nssum).all(1)
    return ok.sum()
for name, fn in (("step_dev", step_dev), ("step_dev_no_h2d", step_dev2), ("step_dev", step_dev)):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(10):
        fn()
    e1.record()
    torch.cuda.synchronize()
    print(name, "ms/step", e0.elapsed_time(e1) / 10)
