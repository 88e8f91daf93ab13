"""Dev experiment (GPU): why an 11 us elementwise kernel follows the space-to-depth stem convolution."""
import os
import sys

import torch
import torch.nn.functional as F

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from rmem_ocu_b200 import ops  # noqa: E402
from rmem_ocu_b200.engine import InferenceResNet  # noqa: E402
from rmem_ocu_b200.model import DeAOTModel, RMemConfig, deterministic_init_  # noqa: E402

dev = torch.device("cuda")
torch.backends.cudnn.benchmark = True
m = DeAOTModel(RMemConfig(latter_mem_len=1))
deterministic_init_(m.state_dict(), 3)
enc = InferenceResNet(m.encoder, m.encoder_projector, dev)
H, W = 480, 864
img = torch.randn(1, 3, H, W, device=dev)
x = torch.empty(1, (H + 1) // 2 + 3, (W + 1) // 2 + 3, 16, device=dev, dtype=torch.bfloat16).permute(0, 3, 1, 2)
ops.pack_image_s2d(img, x)
print("x", x.shape, x.stride(), x.is_contiguous(memory_format=torch.channels_last))
w = enc.conv1_s2d[0]
print("w", w.shape, w.stride(), w.is_contiguous(memory_format=torch.channels_last))
from torch.profiler import profile, ProfilerActivity
for name, fn in [("conv2d", lambda: F.conv2d(x, w, None)),
                 ("conv2d_cl_input", lambda: F.conv2d(x.contiguous(memory_format=torch.channels_last), w, None)),
                 ("cudnn_relu", lambda: torch.cudnn_convolution_relu(x, w, enc.conv1_s2d[1], (1, 1), (0, 0), (1, 1), 1)),
                 ("conv2d_nobench", None)]:
    if fn is None:
        torch.backends.cudnn.benchmark = False
        fn = lambda: F.conv2d(x, w, None)
    for _ in range(3):
        y = fn()
    torch.cuda.synchronize()
    with profile(activities=[ProfilerActivity.CUDA]) as prof:
        y = fn()
        torch.cuda.synchronize()
    print(name, "out stride", y.stride())
    for e in prof.events():
        if e.device_type == torch.autograd.DeviceType.CUDA:
            print("    %-90s %.2f us" % (e.name[:90], e.device_time))
