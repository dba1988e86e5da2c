"""Dev harness: affb200_argsort_f32 against torch.sort(stable=True) (bit-exact order, ties included) and timings of the
sort and of space_filling_cluster at the config-5 token counts.  Run on the GPU box:  python tools/dev_sort.py"""
import sys, torch
sys.path.insert(0, "/root/repo/ml-autofocusformermod_b200")
from affb200 import point_utils as PU
torch.manual_seed(0)
def ev(fn, it=20):
    for _ in range(3): fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(it): fn()
    b.record(); torch.cuda.synchronize()
    return a.elapsed_time(b) / it * 1e3
for B, n in [(1, 1), (1, 33), (3, 257), (2, 2048), (2, 2049), (1, 4100), (1, 6144), (1, 14336), (1, 30720), (256, 3136), (256, 784), (4, 65000)]:
    x = torch.randn(B, n, device="cuda")
    x = (x * 8).round() / 8          # many ties
    for desc in (False, True):
        o = PU.argsort_stable(x, descending=desc)
        ref = x.sort(dim=1, stable=True, descending=desc)[1]
        assert torch.equal(o, ref), (B, n, desc)
    pos = torch.stack([torch.randint(0, 512, (B, n)), torch.randint(0, 256, (B, n))], 2).float().cuda()
    t1 = ev(lambda: PU.argsort_stable(x)); t2 = ev(lambda: x.sort(dim=1, stable=True))
    t3 = ev(lambda: PU.space_filling_cluster(pos, 8, 256, 512)) if n >= 8 else 0
    print("B=%3d n=%6d  argsort_stable %7.1f us   torch.sort %7.1f us   space_filling_cluster %7.1f us" % (B, n, t1, t2, t3))
print("SORT OK")
