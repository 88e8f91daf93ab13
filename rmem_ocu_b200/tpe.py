"""Temporal positional embedding rows per bank position (host logic).

The reference resamples the 4-row ``mem_pos_emb`` table to the T frames currently in the bank
(networks/layers/transformer.py:594-623 AOT, :1140-1169 DeAOT): for T <= 4 a linear,
align_corners=True resample of the first T rows to T rows (an identity); for T > 4: flip ->
legacy 'nearest' resample 4 -> T -> flip.  Both paths pick whole rows, so the bank position t just
selects a row index; the kernels add that row as a per-frame score bias (see csrc/lt_attn.cu).
"""
import numpy as np


def tpe_row_indices(T, rows=4):
    if T < 1:
        raise ValueError("bank length must be >= 1")
    if T <= rows:
        return list(range(T))
    # legacy nearest: src = min(floor(dst * float32(in / out)), in - 1), applied to the flipped table
    scale = np.float32(rows) / np.float32(T)
    out = []
    for t in range(T):
        j = T - 1 - t
        src = min(int(np.floor(np.float32(j) * scale)), rows - 1)
        out.append(rows - 1 - src)
    return out
