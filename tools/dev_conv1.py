"""Dev experiment (GPU): cuDNN formulations of the ResNet stem (7x7 stride-2 conv on 3 channels) at 480x864.
  a) current: input padded to 8 channels, 7x7 s2 p3
  b) space-to-depth: [H/2+3, W/2+3, 16] input (12 used), 4x4 s1 p0 kernel -- same arithmetic
and the max-pool / decoder cast costs."""
import torch
import torch.nn.functional as F

dev = torch.device("cuda")
H, W = 480, 864
BF = torch.bfloat16
CL = torch.channels_last


def timed(fn, iters=20):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        for _ in range(iters):
            fn()
    g.replay()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(5):
        g.replay()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / (5 * iters) * 1e3


torch.backends.cudnn.benchmark = True
torch.manual_seed(0)
img = torch.randn(1, 3, H, W, device=dev)
w = torch.randn(64, 3, 7, 7, device=dev) * 0.05
b = torch.randn(64, device=dev) * 0.1
ref = F.relu(F.conv2d(img, w, b, 2, 3))

# a) current
x8 = torch.zeros(1, 8, H, W, device=dev, dtype=BF).contiguous(memory_format=CL)
x8[:, :3] = img
w8 = torch.zeros(64, 8, 7, 7, device=dev)
w8[:, :3] = w
w8 = w8.to(BF).contiguous(memory_format=CL)
bb = b.to(BF)
fa = lambda: torch.cudnn_convolution_relu(x8, w8, bb, (2, 2), (3, 3), (1, 1), 1)
ya = fa()
print("a) 8ch 7x7 s2: %.2f us, err %.4f" % (timed(fa), (ya.float() - ref).abs().max().item()))

# b) space to depth: out[y,x] = sum_{a,b in -2..1} sum_{dy,dx,c} W[2a+dy+3, 2b+dx+3, c] X[2(y+a)+dy, 2(x+b)+dx, c]
for cpad in (12, 16):
    xs = torch.zeros(1, cpad, H // 2 + 3, W // 2 + 3, device=dev)
    ws = torch.zeros(64, cpad, 4, 4, device=dev)
    for dy in range(2):
        for dx in range(2):
            ch = (dy * 2 + dx) * 3
            xs[:, ch:ch + 3, 2:2 + H // 2, 2:2 + W // 2] = img[:, :, dy::2, dx::2]
            for a in range(-2, 2):
                for bq in range(-2, 2):
                    i, j = 2 * a + dy + 3, 2 * bq + dx + 3
                    if 0 <= i < 7 and 0 <= j < 7:
                        ws[:, ch:ch + 3, a + 2, bq + 2] = w[:, :, i, j]
    chk = F.relu(F.conv2d(xs, ws, b))
    print("   s2d exactness (fp32):", (chk - ref).abs().max().item(), tuple(chk.shape))
    xsb = xs.to(BF).contiguous(memory_format=CL)
    wsb = ws.to(BF).contiguous(memory_format=CL)
    fb = lambda: torch.cudnn_convolution_relu(xsb, wsb, bb, (1, 1), (0, 0), (1, 1), 1)
    yb = fb()
    print("b) s2d %dch 4x4 s1: %.2f us, err %.4f" % (cpad, timed(fb), (yb.float() - ref).abs().max().item()))

fm = lambda: F.max_pool2d(ya, 3, 2, 1)
print("maxpool torch: %.2f us" % timed(fm))
try:
    import sys, os
    sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
    from rmem_ocu_b200 import ops
    if hasattr(ops, "maxpool3s2_nhwc"):
        out = torch.empty(1, 64, H // 4, W // 4, device=dev, dtype=BF).contiguous(memory_format=CL)
        fo = lambda: ops.maxpool3s2_nhwc(ya, out)
        fo()
        print("maxpool rmem: %.2f us, equal %s" % (timed(fo), torch.equal(out, fm())))
except Exception as e:  # dev tool
    print("rmem maxpool unavailable:", e)

# decoder shortcut casts
s4 = torch.randn(1, 256, 120, 216, device=dev, dtype=BF).contiguous(memory_format=CL)
fc = lambda: s4.float()
print("cast 256x120x216 bf16->f32: %.2f us" % timed(fc))
wa = torch.randn(256, 256, 1, 1, device=dev).contiguous(memory_format=CL)
s4f = s4.float()
print("adapter_4x f32 conv: %.2f us" % timed(lambda: F.conv2d(s4f, wa)))
wab = wa.to(BF)
print("adapter_4x bf16 conv: %.2f us" % timed(lambda: F.conv2d(s4, wab)))
