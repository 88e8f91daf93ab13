"""Dev: per-tile SM-clock trace of one CTA of the stream-K attention kernel (library built with -DLT_SK_TRACE=<cta>)."""
import ctypes, sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
from rmem_ocu_b200 import ops, _lib

L, T = int(sys.argv[1]) if len(sys.argv) > 1 else 1620, int(sys.argv[2]) if len(sys.argv) > 2 else 5
mk = lambda *s: torch.randn(*s, device="cuda").to(torch.bfloat16)
q, kb, vb = mk(L, 128), mk(T, L, 128), mk(T, L, 1024)
order = torch.arange(T, device="cuda", dtype=torch.int32)
bias = torch.randn(L, 4, device="cuda"); rows = [min(t, 3) for t in range(T)]
ws = ops.lt_attn_sk_workspace(L, T, "cuda")
for _ in range(3):
    ops.lt_attn_sk_partial(q, kb, vb, order, T, bias, rows, True, ws)
torch.cuda.synchronize()
buf = np.zeros((3, 64, 16), np.int64)
h = ctypes.CDLL(_lib.LIB_PATH)
assert h.rmem_dev_sk_trace(buf.ctypes.data_as(ctypes.c_void_p)) == 0
t0 = buf[buf > 0].min()
n = int((buf[1, :, 6] > 0).sum())
print("tiles traced:", n)
print("QK issuer: wait k_full | issue+commit || PV issuer: wait v_full(+o_drained) | wait p_full | issue+commit | period")
prev = None
for i in range(n):
    m = buf[1, i] - t0
    per = (m[6] - prev) if prev is not None else 0
    prev = m[6]
    print("tile %2d  qk@%6d kwait %5d issue %4d || pv@%6d vwait %5d pwait %5d issue %4d | period %5d" %
          (i, m[0], m[1] - m[0], m[2] - m[1], m[3], m[4] - m[3], m[5] - m[4], m[6] - m[5], per))
print("softmax (first softmax warp, lane 0): wait s_full | ld+max | pair bar | exp | st+arrive | (epilogue)")
for i in range(n):
    s = buf[2, i] - t0
    print("tile %2d  @%6d swait %5d ldmax %5d bar %4d exp %5d st %4d epi %5d" %
          (i, s[0], s[1] - s[0], s[2] - s[1], s[3] - s[2], s[4] - s[3], s[5] - s[4], max(0, s[6] - s[5]) if buf[2, i, 6] > 0 else 0))
print("producer: wait k_empty | wait v_empty")
for i in range(n):
    p = buf[0, i] - t0
    print("tile %2d  @%6d kempty %5d vempty %5d" % (i, p[0], p[1] - p[0], p[2] - p[1]))
