import sys, torch
sys.path.insert(0, "ml-autofocusformermod_b200")
from affb200.norm_ops import ConvStemFunction
x = torch.randn(256, 3, 224, 224, device="cuda").bfloat16().contiguous(memory_format=torch.channels_last)
conv = torch.nn.Conv2d(3, 32, 3, 2, 1).cuda()
wl = conv.weight.detach().bfloat16().contiguous(memory_format=torch.channels_last)
def t(f, n=20):
    for _ in range(3): f()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(True), torch.cuda.Event(True)
    a.record()
    for _ in range(n): f()
    b.record(); torch.cuda.synchronize()
    return a.elapsed_time(b) / n * 1e3
with torch.no_grad():
    print("ours  %.1f us" % t(lambda: ConvStemFunction.apply(x, conv.weight)))
    print("cudnn %.1f us" % t(lambda: torch.nn.functional.conv2d(x, wl, None, 2, 1)))
g = torch.randn(256, 32, 112, 112, device="cuda").bfloat16().contiguous(memory_format=torch.channels_last)
print("cudnn wgrad %.1f us" % t(lambda: torch.ops.aten.convolution_backward(g, x, wl, None, [2, 2], [1, 1], [1, 1], False, [0, 0], 1, [False, True, False])))
for cpad in (4, 8):
    xp = torch.zeros(256, cpad, 224, 224, device="cuda", dtype=torch.bfloat16).contiguous(memory_format=torch.channels_last)
    xp[:, :3] = x
    wp = torch.zeros(32, cpad, 3, 3, device="cuda", dtype=torch.bfloat16).contiguous(memory_format=torch.channels_last)
    print("cudnn wgrad, input padded to %d channels %.1f us" % (cpad, t(lambda: torch.ops.aten.convolution_backward(
        g, xp, wp, None, [2, 2], [1, 1], [1, 1], False, [0, 0], 1, [False, True, False]))))
    print("   pad copy %.1f us" % t(lambda: xp[:, :3].copy_(x)))
