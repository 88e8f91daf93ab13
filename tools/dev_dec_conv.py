"""The decoder's four 3x3 convolutions at 480p (30x54 .. 120x216) through cuDNN: TF32 on fp32 tensors vs fp16."""
import torch, torch.nn.functional as F
torch.backends.cudnn.benchmark = True
shapes = [("conv_in", 768, 256, 30, 54), ("conv_16x", 256, 256, 30, 54), ("conv_8x", 256, 128, 60, 108), ("conv_4x", 128, 128, 120, 216)]
for dt in (torch.float32, torch.float16, torch.bfloat16):
    tot = 0
    for name, ci, co, h, w in shapes:
        x = torch.randn(1, ci, h, w, device="cuda", dtype=dt).contiguous(memory_format=torch.channels_last)
        wt = torch.randn(co, ci, 3, 3, device="cuda", dtype=dt).contiguous(memory_format=torch.channels_last)
        for _ in range(3): F.conv2d(x, wt, None, 1, 1)
        torch.cuda.synchronize()
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g):
            for _ in range(10): y = F.conv2d(x, wt, None, 1, 1)
        g.replay(); torch.cuda.synchronize()
        a = torch.cuda.Event(enable_timing=True); b = torch.cuda.Event(enable_timing=True)
        a.record(); [g.replay() for _ in range(10)]; b.record(); torch.cuda.synchronize()
        t = a.elapsed_time(b) * 1e3 / 100
        tot += t
        print(dt, name, "%.1f us" % t)
    print(dt, "total %.1f us" % tot)
