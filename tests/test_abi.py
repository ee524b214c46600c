"""CPU suite: the C-ABI library loads and exports every symbol include/bt2gpu.h declares;
without a GPU every compute entry point fails loudly (no CPU fallback)."""
import ctypes as C
import os
import re

import pytest

import helpers as H
from bt2gpu import lib, structs


def _declared_symbols():
    txt = open(os.path.join(H.ROOT, "include", "bt2gpu.h")).read()
    return sorted(set(re.findall(r"\b(bt2gpu_[a-z0-9_]+)\s*\(", txt)))


def test_library_exports_every_declared_symbol():
    l = lib.load()
    names = _declared_symbols()
    assert len(names) >= 14
    for n in names:
        assert hasattr(l, n), n
    assert sorted(lib.SYMBOLS) == names
    assert l.bt2gpu_abi_version() == 2


def test_struct_sizes_match_header():
    # sizes are asserted at import of bt2gpu.structs; spot-check the offsets the kernels rely on
    assert structs.dp_job.fields["refl"][1] == 24
    assert structs.dp_aln.fields["refoff"][1] == 16


def test_no_cpu_fallback():
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    with pytest.raises(lib.Bt2GpuError):
        lib.Context(0)
