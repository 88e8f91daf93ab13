"""Bring-up check for the long-term attention kernels on a B200 (dev tool, run under gpurun).

Staged so that a failure localises the broken primitive:
  A  Q = 0           -> P uniform; exercises TMA(V) + P smem tile + PV MMA (MN-major V descriptor)
  B  V = const       -> checks m/l (log-sum-exp) of QK^T: TMA(Q,K) + K-major descriptors
  C  random          -> full kernel incl. lazy rescale, masks, ragged L, several frames / chunks
"""
import math
import sys
import time

import torch

sys.path.insert(0, ".")
from rmem_ocu_b200 import _lib  # noqa: E402


def ptr(t):
    return t.data_ptr() if t is not None else None


def run(q, kb, vb, order, T, cpf, tpe_bias=None, tpe_rows=None, gate=None, want_mass=True):
    lib = _lib.lib()
    L = q.shape[0]
    n_slots = kb.shape[0]
    nch = T * cpf
    po = torch.empty(nch, L, 1024, device="cuda", dtype=torch.float32)
    pm = torch.empty(nch, L, device="cuda", dtype=torch.float32)
    pl = torch.empty(nch, L, device="cuda", dtype=torch.float32)
    out = torch.empty(L, 1024, device="cuda", dtype=torch.bfloat16)
    mass = torch.empty(L, T, device="cuda", dtype=torch.float32) if want_mass else None
    st = torch.cuda.current_stream().cuda_stream
    _lib.check(lib.rmem_lt_attn_deaot_partial(ptr(q), ptr(kb), ptr(vb), ptr(order), T, L, n_slots, cpf,
                                              ptr(po), ptr(pm), ptr(pl), st), "partial")
    import ctypes
    rows = (ctypes.c_int32 * T)(*(tpe_rows if tpe_rows is not None else [0] * T))
    _lib.check(lib.rmem_lt_attn_deaot_combine(ptr(po), ptr(pm), ptr(pl), ptr(tpe_bias),
                                              ctypes.cast(rows, ctypes.c_void_p), T, L, cpf, ptr(gate), ptr(out),
                                              ptr(mass), st), "combine")
    torch.cuda.synchronize()
    return out, mass, (po, pm, pl)


def ref(q, kb, vb, order, T, tpe_bias_nat=None, tpe_rows=None, gate=None):
    """fp32 reference on the bf16-rounded inputs. tpe_bias_nat: [L,4] additive score bias (natural units)."""
    L = q.shape[0]
    qf = q.float()
    scores = []
    for t in range(T):
        k = kb[order[t]].float()
        s = qf @ k.t() / math.sqrt(128.0)
        if tpe_bias_nat is not None:
            s = s + tpe_bias_nat[:, tpe_rows[t]][:, None]
        scores.append(s)
    s = torch.cat(scores, 1)
    a = torch.softmax(s, -1)
    v = torch.cat([vb[order[t]].float() for t in range(T)], 0)
    o = a @ v
    if gate is not None:
        o = o * gate.float()
    mass = a.view(L, T, L).sum(-1)
    return o, mass


def report(name, out, o_ref, mass, m_ref):
    err = (out.float() - o_ref).abs()
    rel = err.norm() / o_ref.norm().clamp_min(1e-12)
    me = (mass - m_ref).abs().max().item() if mass is not None else float("nan")
    print(f"[{name}] out max_abs_err={err.max().item():.4e} rel_fro={rel.item():.4e} "
          f"mass max_abs_err={me:.3e} finite={bool(torch.isfinite(out.float()).all())}")
    if rel.item() > 2e-2 or not math.isfinite(rel.item()):
        # show where it goes wrong (row / column structure)
        e = err.view(err.shape[0], -1)
        rows = e.max(1).values
        cols = e.max(0).values
        print("   worst rows:", rows.topk(min(8, rows.numel())).indices.tolist())
        print("   err by 64-col block:", [round(x, 3) for x in cols.view(-1, 64).max(1).values.tolist()])
        print("   err by 128-row block:", [round(x, 3) for x in
                                          torch.nn.functional.pad(rows, (0, (-rows.numel()) % 128)).view(-1, 128).max(1).values.tolist()])
        print("   out[0,:8]=", out[0, :8].float().tolist(), " ref[0,:8]=", o_ref[0, :8].tolist())
        return False
    return True


def main():
    torch.manual_seed(0)
    dev = "cuda"
    ok = True
    # ---- A: PV path only
    for (L, T, cpf) in [(128, 1, 1), (200, 2, 1)]:
        n_slots = T + 1
        q = torch.zeros(L, 128, device=dev, dtype=torch.bfloat16)
        kb = torch.randn(n_slots, L, 128, device=dev).to(torch.bfloat16)
        vb = torch.randn(n_slots, L, 1024, device=dev).to(torch.bfloat16)
        order = torch.arange(T, device=dev, dtype=torch.int32)
        out, mass, _ = run(q, kb, vb, order, T, cpf)
        o_ref, m_ref = ref(q, kb, vb, order.tolist(), T)
        ok &= report(f"A PV-only L={L} T={T}", out, o_ref, mass, m_ref)
    # ---- B: QK path only (V const) -> compare log-sum-exp
    L, T, cpf = 256, 1, 1
    q = torch.randn(L, 128, device=dev).to(torch.bfloat16)
    kb = torch.randn(2, L, 128, device=dev).to(torch.bfloat16)
    vb = torch.ones(2, L, 1024, device=dev, dtype=torch.bfloat16)
    order = torch.zeros(1, device=dev, dtype=torch.int32)
    out, mass, (po, pm, pl) = run(q, kb, vb, order, T, cpf)
    s = q.float() @ kb[0].float().t() / math.sqrt(128.0)
    lse_ref = torch.logsumexp(s, -1) / math.log(2.0)
    lse = pm[0] + torch.log2(pl[0])
    print(f"[B QK-only] lse max_abs_err={(lse - lse_ref).abs().max().item():.4e}  out==1 err={(out.float() - 1).abs().max().item():.3e}")
    ok &= (lse - lse_ref).abs().max().item() < 5e-2
    # ---- C: full random, ragged, several frames, permuted slots, TPE bias, gate
    for (L, T, cpf, scale) in [(1705, 5, 1, 1.0), (1620, 3, 2, 3.0), (333, 9, 1, 6.0), (1705, 1, 3, 1.0)]:
        n_slots = T + 1
        q = (torch.randn(L, 128, device=dev) * scale).to(torch.bfloat16)
        kb = torch.randn(n_slots, L, 128, device=dev).to(torch.bfloat16)
        vb = torch.randn(n_slots, L, 1024, device=dev).to(torch.bfloat16)
        gate = torch.randn(L, 1024, device=dev).to(torch.bfloat16)
        perm = torch.randperm(n_slots)[:T]
        order = perm.to(device=dev, dtype=torch.int32)
        rows = [min(t, 3) for t in range(T)]
        bias_nat = torch.randn(L, 4, device=dev) * 0.5
        bias_log2 = (bias_nat * math.log2(math.e)).contiguous()
        out, mass, _ = run(q, kb, vb, order, T, cpf, tpe_bias=bias_log2, tpe_rows=rows, gate=gate)
        o_ref, m_ref = ref(q, kb, vb, perm.tolist(), T, bias_nat, rows, gate)
        ok &= report(f"C full L={L} T={T} cpf={cpf} qscale={scale}", out, o_ref, mass, m_ref)
    # ---- timing of the headline shape
    L, T, cpf = 1705, 5, 1
    q = torch.randn(L, 128, device=dev).to(torch.bfloat16)
    kb = torch.randn(T + 1, L, 128, device=dev).to(torch.bfloat16)
    vb = torch.randn(T + 1, L, 1024, device=dev).to(torch.bfloat16)
    order = torch.arange(T, device=dev, dtype=torch.int32)
    for _ in range(3):
        run(q, kb, vb, order, T, cpf, want_mass=False)
    lib = _lib.lib()
    nch = T * cpf
    po = torch.empty(nch, L, 1024, device=dev, dtype=torch.float32)
    pm = torch.empty(nch, L, device=dev, dtype=torch.float32)
    pl = torch.empty(nch, L, device=dev, dtype=torch.float32)
    st = torch.cuda.current_stream().cuda_stream
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(20):
        lib.rmem_lt_attn_deaot_partial(ptr(q), ptr(kb), ptr(vb), ptr(order), T, L, T + 1, cpf, ptr(po), ptr(pm), ptr(pl), st)
    e1.record()
    torch.cuda.synchronize()
    us = e0.elapsed_time(e1) / 20 * 1e3
    flops = 2.0 * L * (T * L) * (128 + 1024)
    print(f"[time] partial kernel L={L} T={T}: {us:.1f} us  -> {flops / us / 1e6:.1f} TFLOP/s useful")
    print("ALL OK" if ok else "SOME FAILED")
    return 0 if ok else 1


if __name__ == "__main__":
    sys.exit(main())
