"""The C-ABI shared library builds for sm_100a without a GPU, loads, and exports every entry point
include/rmem_b200.h declares; the Python side fails loudly (no CPU fallback) without CUDA."""
import ctypes
import re

import pytest
import torch


def test_library_exports_every_declared_symbol(lib_built):
    lib = lib_built.lib()
    declared = lib_built.declared_symbols()
    assert len(declared) >= 15
    for sym in declared:
        assert hasattr(lib, sym), sym
    # every declared entry point has a ctypes signature (so a renamed symbol cannot slip through)
    assert sorted(lib_built._SIGNATURES) == sorted(declared)


def test_abi_version_and_error_string(lib_built):
    lib = lib_built.lib()
    with open(lib_built.HEADER_PATH) as f:
        ver = int(re.search(r"#define RMEM_ABI_VERSION (\d+)", f.read()).group(1))
    assert lib.rmem_abi_version() == ver
    assert isinstance(lib.rmem_last_error(), (bytes, type(None)))


def test_header_constants_match_python(lib_built):
    with open(lib_built.HEADER_PATH) as f:
        text = f.read()
    get = lambda n: int(re.search(r"#define %s (\d+)" % n, text).group(1))
    assert get("RMEM_MAX_BANK") == lib_built.MAX_BANK
    assert get("RMEM_MAX_CHUNKS_PER_FRAME") == lib_built.MAX_CHUNKS_PER_FRAME
    assert lib_built.BANK_STATE_WORDS == 5 * lib_built.MAX_BANK + 8
    assert ctypes.sizeof(lib_built.GemmSegment) == 32


def test_sass_has_blackwell_tensor_and_tma_instructions(lib_built):
    """The attention kernels really are tcgen05 / TMA kernels (B200_PROFILING.md: UTC*MMA, UTMALDG)."""
    import subprocess
    out = subprocess.run(["/usr/local/cuda/bin/cuobjdump", "-sass", lib_built.LIB_PATH], capture_output=True,
                         text=True).stdout
    assert "UTCHMMA" in out or "UTCMMA" in out
    assert "UTMALDG" in out
    assert "LDTM" in out


@pytest.mark.skipif(torch.cuda.is_available(), reason="checks the no-GPU failure mode")
def test_engine_refuses_to_run_without_cuda(lib_built):
    from rmem_ocu_b200 import ops
    from rmem_ocu_b200.engine import DeAOTEngine
    from rmem_ocu_b200.model import DeAOTModel
    with pytest.raises(lib_built.RmemError):
        DeAOTEngine(DeAOTModel())
    with pytest.raises(lib_built.RmemError):
        ops.layernorm(torch.zeros(4, 256), torch.ones(256), torch.zeros(256), torch.zeros(4, 256))
