"""The engines accept the reference's OWN nn.Module (networks/engines/__init__.py:5-21 hands build_engine the
real AOT / DeAOT object): the weight repack (`engine.prepare`) runs on a model constructed by /root/reference
code, on CPU, without mutating it, and the folded-BN encoder it builds reproduces the reference encoder.
Build container only (needs /root/reference)."""
import pytest
import torch

from oracle import ref_loader as RL

pytestmark = [pytest.mark.reference,
              pytest.mark.skipif(not RL.available(), reason="reference not present at /root/reference")]


@pytest.mark.parametrize("name", ["r50_deaotl", "r50_aotl"])
def test_prepare_accepts_reference_built_module(name, monkeypatch):
    from rmem_ocu_b200 import engine as E
    from rmem_ocu_b200.model import deterministic_init_
    monkeypatch.setenv("RMEM_ENCODER_DTYPE", "fp32")
    model = RL.build_model(RL.make_cfg(name, 1, 4))
    deterministic_init_(model.state_dict(), 3)
    before = {k: v.clone() for k, v in model.state_dict().items()}
    assert not hasattr(model.encoder, "fold_bn_")                  # this is the reference's ResNet, not ours
    prep = E.prepare(model, torch.device("cpu"))
    assert prep.deaot == (name == "r50_deaotl")
    assert prep.num_layers == 3
    after = model.state_dict()
    assert all(torch.equal(before[k], after[k]) for k in before), "prepare() must not mutate the caller's model"
    # the folded encoder + projector == the reference's encoder + projector (fp32, CPU)
    img = torch.randn(1, 3, 96, 128, generator=torch.Generator().manual_seed(5))
    with torch.no_grad():
        want = list(model.encoder(img))
        want[-1] = model.encoder_projector(want[-1])
        got = prep.encoder(img)
    assert len(got) == len(want) == 4
    for a, b in zip(got, want):
        assert a.shape == b.shape
        assert (a - b).abs().max().item() <= 2e-4 * max(1.0, b.abs().max().item())


def test_object_group_split_and_merge_match_reference():
    """separate_mask / soft_logit_aggregation of AOTInferEngine (aot_engine.py:604-628, 650-673) are pure tensor
    functions: the product's and the oracle's restatements must reproduce the reference's outputs exactly."""
    import types
    from oracle import restate as R
    from rmem_ocu_b200 import engine as E
    ref = RL.load().aot_engine.AOTInferEngine
    stub = types.SimpleNamespace(aot_engines=[None, None, None], max_aot_obj_num=10)
    g = torch.Generator().manual_seed(9)
    mask = torch.randint(0, 26, (1, 1, 37, 53), generator=g).float()
    logits = [torch.randn(1, 11, 37, 53, generator=g) * 4 for _ in range(3)]
    want_masks = ref.separate_mask(stub, mask)
    want_logit = ref.soft_logit_aggregation(stub, logits)
    for impl in (E.DeAOTInferEngine, R.InferEngine):
        got = impl.separate_mask(stub, mask)
        assert len(got) == 3 and all(torch.equal(a, b) for a, b in zip(got, want_masks)), impl
        assert torch.equal(impl.soft_logit_aggregation(stub, logits), want_logit), impl
