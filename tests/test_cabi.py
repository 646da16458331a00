"""CPU-side checks of the boundary: the library loads and exports every symbol include/hmm_b200.h declares."""
import ctypes
import os
import re

from hmm_layer_b200 import _lib

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_symbols():
    hdr = open(os.path.join(ROOT, "include", "hmm_b200.h")).read()
    hdr = re.sub(r"/\*.*?\*/", "", hdr, flags=re.S)
    return sorted(set(re.findall(r"\b(hmm_b200_\w+)\s*\(", hdr)))


def test_library_exports_every_declared_symbol():
    lib = ctypes.CDLL(_lib.LIB_PATH)
    syms = declared_symbols()
    assert len(syms) >= 6
    for s in syms:
        assert hasattr(lib, s), f"{s} declared in include/hmm_b200.h but not exported"
    assert sorted(_lib.EXPORTS) == syms, "hmm_layer_b200/_lib.py EXPORTS out of sync with the header"


def test_abi_version_and_workspace_query_need_no_gpu():
    lib = _lib.load()
    assert lib.hmm_b200_abi_version() == 1
    assert lib.hmm_b200_genepred_workspace_size(1, 32, 9999) > 0
    assert lib.hmm_b200_genepred_workspace_size(0, 32, 9999) == 0
    assert lib.hmm_b200_set_tuning(64, 0, 0) != 0      # chunk length above the shared-memory bound rejected
    assert lib.hmm_b200_set_tuning(32, 100, 0) != 0    # finalize threads must be a multiple of 32
    assert lib.hmm_b200_set_tuning(32, 256, 0) == 0
    assert lib.hmm_b200_set_tuning(0, 0, 0) == 0
