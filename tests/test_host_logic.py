"""Host-side logic that needs no GPU: temporal-PE row selection, sine position embedding, split-KV grid
choice, model-family dispatch -- each against the oracle restatement (which is pinned to the reference)."""
import pytest
import torch

from oracle import restate as R


@pytest.mark.parametrize("T", list(range(1, 17)))
def test_tpe_rows_match_oracle(T):
    from rmem_ocu_b200.tpe import tpe_row_indices
    assert tpe_row_indices(T) == R.tpe_row_indices(T)


@pytest.mark.parametrize("h,w", [(30, 54), (31, 55), (6, 7), (45, 80)])
def test_sine_position_embedding_matches_oracle(h, w):
    from rmem_ocu_b200.engine_aot import position_embedding_sine
    ours = position_embedding_sine(h, w)                                           # [L, 256]
    ref = R.position_embedding_sine(h, w).view(1, -1, h * w).permute(2, 0, 1)[:, 0]
    assert ours.shape == ref.shape
    assert torch.equal(ours, ref)


def test_chunk_choice_fills_the_gpu_and_respects_limits():
    from rmem_ocu_b200 import _lib
    from rmem_ocu_b200.engine import _chunks_per_frame
    from rmem_ocu_b200.engine_aot import mha_chunks_per_frame
    for fn in (_chunks_per_frame, mha_chunks_per_frame):
        for L in (42, 1620, 1705, 3600):
            for T in range(1, _lib.MAX_BANK + 1):
                cpf = fn(T, L)
                assert 1 <= cpf <= _lib.MAX_CHUNKS_PER_FRAME
        assert fn(1, 1620) > 1            # a single memory frame must still be split across the SMs


def test_model_family_dispatch():
    from rmem_ocu_b200.engine import is_deaot_model
    from rmem_ocu_b200.model import AOTModel, DeAOTModel
    assert is_deaot_model(DeAOTModel())
    assert not is_deaot_model(AOTModel())


def test_aot_weight_repack_is_algebraically_the_reference_block():
    """The concatenated / block weights of PreparedAOT reproduce the separate Linears: checked on CPU in fp32
    by evaluating both forms (no kernels involved)."""
    from tests import _synth
    _, sd = _synth.models(3, vos="aot")
    p = "LSTT.layers.0."
    x, e = torch.randn(5, 256), torch.randn(5, 256)
    wv, bv = sd[p + "linear_V.weight"], sd[p + "linear_V.bias"]
    direct = torch.nn.functional.linear(x + e, wv, bv)
    packed = torch.nn.functional.linear(torch.cat([x, e], 1), torch.cat([wv, wv], 1), bv)
    assert torch.allclose(direct, packed, atol=1e-5)
    lt, st = torch.randn(5, 256), torch.randn(5, 256)
    wl, ws = sd[p + "long_term_attn.projection.weight"], sd[p + "short_term_attn.projection.weight"]
    bl, bs = sd[p + "long_term_attn.projection.bias"], sd[p + "short_term_attn.projection.bias"]
    z = torch.zeros(256, 256)
    wbig = torch.cat([torch.cat([wl, ws], 1), torch.cat([z, ws], 1)], 0)
    out = torch.nn.functional.linear(torch.cat([lt, st], 1), wbig, torch.cat([bl + bs, bs]))
    tgt2 = torch.nn.functional.linear(lt, wl, bl)
    tgt3 = torch.nn.functional.linear(st, ws, bs)
    assert torch.allclose(out[:, :256], tgt2 + tgt3, atol=1e-5)
    assert torch.allclose(out[:, 256:], tgt3, atol=1e-5)
