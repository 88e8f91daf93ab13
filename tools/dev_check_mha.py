"""Bring-up helper for csrc/mha.cu: launches one small case asynchronously and, if it does not finish,
prints the progress markers the kernel left in mapped host memory."""
import ctypes, math, sys, time
import torch
sys.path.insert(0, ".")
from rmem_ocu_b200 import _lib, ops

L, T, cpf = int(sys.argv[1]), int(sys.argv[2]), int(sys.argv[3])
lib = ctypes.CDLL(_lib.LIB_PATH)
dbg = torch.zeros(8 * 4096, dtype=torch.int32).pin_memory()
if hasattr(lib, "rmem_debug_set_buffer"):
    lib.rmem_debug_set_buffer.argtypes = [ctypes.c_void_p]
    lib.rmem_debug_set_buffer(dbg.data_ptr())
torch.manual_seed(0)
q = torch.randn(L, 256, device="cuda").bfloat16()
kb = torch.randn(T, L, 256, device="cuda").bfloat16()
vb = torch.randn(T, L, 256, device="cuda").bfloat16()
order = torch.arange(T, device="cuda", dtype=torch.int32)
nch = T * cpf
po = torch.zeros(nch, L, 256, device="cuda"); pm = torch.zeros(nch, 8, L, device="cuda"); pl = torch.zeros(nch, 8, L, device="cuda")
out = torch.zeros(L, 256, device="cuda", dtype=torch.bfloat16)
mass = torch.zeros(L, T, device="cuda")
torch.cuda.synchronize()
ev = torch.cuda.Event()
ops.mha_partial(q, kb, vb, order, T, cpf, po, pm, pl)
ev.record()
t0 = time.time()
while not ev.query() and time.time() - t0 < 8:
    time.sleep(0.05)
n_cta = -(-L // 128) * 4 * nch
print("partial done:", ev.query(), flush=True)
d = dbg[:8 * n_cta].view(n_cta, 8)
bad = [(i, d[i].tolist()) for i in range(n_cta) if d[i, 7] == 0]
print("unfinished CTAs (x + 4-pair*qtiles.. linear: ((z*4+y)*qt+x)):", len(bad), bad[:12], flush=True)
if not ev.query():
    sys.exit(3)
ops.mha_combine(po, pm, pl, None, None, T, L, cpf, out, 0, mass)
torch.cuda.synchronize()
qh = q.float().view(L, 8, 32).permute(1, 0, 2)
s = torch.cat([qh @ kb[t].float().view(L, 8, 32).permute(1, 2, 0) / math.sqrt(32.0) for t in range(T)], 2)
a = torch.softmax(s, -1)
v = torch.cat([vb[t].float() for t in range(T)], 0).view(T * L, 8, 32).permute(1, 0, 2)
ref = (a @ v).permute(1, 0, 2).reshape(L, 256)
print("rel err", ((out.float() - ref).norm() / ref.norm()).item(), "mass err",
      (mass - a.mean(0).view(L, T, L).sum(-1)).abs().max().item())
