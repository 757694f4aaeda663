import torch, sys, os
sys.path.insert(0, os.getcwd())
from onepose_st_b200.nn_common import ResNetFPN_8_2
from torch.profiler import profile, ProfilerActivity
m = ResNetFPN_8_2({"block_type": "BasicBlock", "initial_dim": 128, "block_dims": [128, 196, 256]}).cuda().eval()
x = torch.rand(8, 1, 512, 512, device="cuda")
with torch.no_grad():
    for _ in range(3): m(x)
    torch.cuda.synchronize()
    with profile(activities=[ProfilerActivity.CPU, ProfilerActivity.CUDA], record_shapes=True) as prof:
        m(x)
        torch.cuda.synchronize()
evs = [e for e in prof.events() if e.device_type == torch.autograd.DeviceType.CUDA]
# map kernels to the enclosing CPU op by time correlation id
rows = []
for e in prof.events():
    if e.device_type == torch.autograd.DeviceType.CPU and e.kernels:
        for k in e.kernels:
            rows.append((k.duration, e.name, str(e.input_shapes)[:90], k.name[:70]))
seen = set()
for d, n, sh, kn in sorted(rows, key=lambda r: -r[0])[:60]:
    key = (n, sh, kn)
    print(f"{d:9.1f} us  {n:40s} {sh:90s} {kn}")
