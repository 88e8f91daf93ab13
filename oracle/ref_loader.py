"""TEST INFRASTRUCTURE ONLY -- loader for the *live* reference (Bardli/RMem_ocu).

This module imports the unmodified reference from ``/root/reference/aot_plus``.
It only works in the build container (the GPU box has no ``/root/reference``);
it is used (a) by ``oracle/make_golden.py`` to generate the committed fixtures
under ``tests/golden/`` and (b) by ``tests/test_oracle_vs_reference.py`` to pin
the CPU restatement in ``oracle/restate.py`` against the reference itself.
Nothing under ``rmem_ocu_b200/`` may import this file.

Harness-side shims (none of them touches /root/reference):
  * ``timm`` is not installed: ``aot.py:9`` / ``deaot.py:7`` import
    ``timm.models.layers.trunc_normal_``; the swin encoder chain also imports
    ``DropPath``/``to_2tuple``.  We register stub modules.
  * ``matplotlib`` is not installed: ``layers/position.py:2`` imports pyplot.
  * ``configs/pre_*.py`` are import-broken; we build
    ``configs.default.EngineConfig`` directly and set the two attributes
    ``pre_vost.py:15-17`` would set.
  * ``AOTEngine.assign_identity`` (``aot_engine.py:209-213``) hard-codes a cuda
    device for the default ignore mask; we pass a CPU zeros mask explicitly.
"""
import contextlib
import io
import os
import sys
import types

REFERENCE_ROOT = os.environ.get("RMEM_REFERENCE_ROOT", "/root/reference")
_AOT_PLUS = os.path.join(REFERENCE_ROOT, "aot_plus")


def available() -> bool:
    return os.path.isdir(os.path.join(_AOT_PLUS, "networks"))


def _install_stubs():
    import torch

    if "timm" not in sys.modules:
        timm = types.ModuleType("timm")
        models = types.ModuleType("timm.models")
        layers = types.ModuleType("timm.models.layers")

        def trunc_normal_(tensor, mean=0., std=1., a=-2., b=2.):
            return torch.nn.init.trunc_normal_(tensor, mean=mean, std=std, a=a, b=b)

        class DropPath(torch.nn.Module):
            def __init__(self, drop_prob=None):
                super().__init__()

            def forward(self, x):
                return x

        def to_2tuple(x):
            return x if isinstance(x, tuple) else (x, x)

        layers.trunc_normal_ = trunc_normal_
        layers.DropPath = DropPath
        layers.to_2tuple = to_2tuple
        models.layers = layers
        timm.models = models
        sys.modules["timm"] = timm
        sys.modules["timm.models"] = models
        sys.modules["timm.models.layers"] = layers
    if "matplotlib" not in sys.modules:
        mpl = types.ModuleType("matplotlib")
        plt = types.ModuleType("matplotlib.pyplot")
        mpl.pyplot = plt
        sys.modules["matplotlib"] = mpl
        sys.modules["matplotlib.pyplot"] = plt


_loaded = {}


def load():
    """Import the reference packages; returns a namespace of the modules we use."""
    if _loaded:
        return _loaded["ns"]
    if not available():
        raise RuntimeError("reference not present at %s" % REFERENCE_ROOT)
    _install_stubs()
    if _AOT_PLUS not in sys.path:
        sys.path.insert(0, _AOT_PLUS)
    import torch

    rng_state = torch.get_rng_state()  # transformer.py:126-131 consumes RNG at import
    import configs.default as cfg_default
    import networks.engines as engines
    import networks.engines.aot_engine as aot_engine
    import networks.layers.attention as attention
    import networks.layers.basic as basic
    import networks.layers.transformer as transformer
    import networks.models as models
    torch.set_rng_state(rng_state)

    # assign_identity device shim
    if not getattr(aot_engine.AOTEngine, "_rmem_shimmed", False):
        orig = aot_engine.AOTEngine.assign_identity

        def assign_identity(self, one_hot_mask, ignore_mask=None):
            if ignore_mask is None:
                ignore_mask = torch.zeros(one_hot_mask.shape[0], 1, one_hot_mask.shape[2],
                                          one_hot_mask.shape[3], device=one_hot_mask.device)
            return orig(self, one_hot_mask, ignore_mask)

        aot_engine.AOTEngine.assign_identity = assign_identity
        aot_engine.AOTEngine._rmem_shimmed = True

    ns = types.SimpleNamespace(cfg_default=cfg_default, engines=engines, aot_engine=aot_engine,
                               attention=attention, basic=basic, transformer=transformer,
                               models=models)
    _loaded["ns"] = ns
    return ns


def make_cfg(model="r50_deaotl", former=1, latter=4):
    ns = load()
    cfg = ns.cfg_default.EngineConfig("oracle", model)
    cfg.MODEL_LINEAR_Q = False          # configs/pre_vost.py:15-17
    cfg.MODEL_IGNORE_TOKEN = True
    cfg.FORMER_MEM_LEN = former
    cfg.LATTER_MEM_LEN = latter
    return cfg


def build_model(cfg, seed=0):
    import torch
    ns = load()
    torch.manual_seed(seed)
    model = ns.models.build_vos_model(cfg.MODEL_VOS, cfg)
    model.eval()
    return model


def build_engine(model, cfg, long_term_mem_gap=5):
    ns = load()
    eng = ns.engines.build_engine(cfg.MODEL_ENGINE, phase="eval", aot_model=model, gpu_id=0,
                                  long_term_mem_gap=long_term_mem_gap)
    eng.eval()
    return eng


@contextlib.contextmanager
def quiet():
    """The reference prints unconditionally inside restrict_long_memories."""
    buf = io.StringIO()
    with contextlib.redirect_stdout(buf):
        yield buf
