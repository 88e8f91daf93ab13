"""Live check (build container only): re-runs the UNMODIFIED reference from /root/reference and
confirms the committed fixtures under tests/golden/ are what it produces today, and that the product's
parameter containers carry exactly the reference's state_dict keys.  Skipped where the reference is
absent (the GPU box)."""
import os

import numpy as np
import pytest

from oracle import ref_loader as RL

pytestmark = [pytest.mark.reference,
              pytest.mark.skipif(not RL.available(), reason="reference not present at /root/reference")]

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


@pytest.mark.parametrize("name", ["evict_aot", "evict_deaot", "ops_aot", "clip_deaot_small", "clip_aot_small"])
def test_fixture_is_current(name):
    from oracle import make_golden as MG
    fresh = MG.GENERATORS[name]()
    stored = np.load(os.path.join(GOLDEN, name + ".npz"))
    assert sorted(fresh.keys()) == sorted(stored.files)
    for k in stored.files:
        a, b = np.asarray(fresh[k]), stored[k]
        if a.dtype.kind in "fc":
            assert np.allclose(a, b, rtol=0, atol=1e-6 * max(1.0, float(np.abs(b).max()))), (name, k)
        else:
            assert np.array_equal(a, b), (name, k)


@pytest.mark.parametrize("name,vos", [("r50_deaotl", "deaot"), ("r50_aotl", "aot")])
def test_state_dict_keys_match_reference(name, vos):
    from rmem_ocu_b200.model import AOTModel, DeAOTModel, RMemConfig
    ref = RL.build_model(RL.make_cfg(name, 1, 4)).state_dict()
    mine = (DeAOTModel if vos == "deaot" else AOTModel)(RMemConfig(model_vos=vos)).state_dict()
    assert {k: tuple(v.shape) for k, v in ref.items()} == {k: tuple(v.shape) for k, v in mine.items()}
