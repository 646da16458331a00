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


def test_no_alloc_test_file_is_collected_last():
    """tests/test_zz_no_alloc.py must stay the last file pytest collects (a failure there must
    not hide other tests); pytest collects a directory's files in name order."""
    import glob
    import os
    files = sorted(os.path.basename(f) for f in glob.glob(os.path.join(os.path.dirname(__file__), "test_*.py")))
    assert files[-1] == "test_zz_no_alloc.py", files[-3:]


def test_header_states_the_static_scratch_limit():
    import os
    import re
    h = open(os.path.join(os.path.dirname(__file__), "..", "include", "hmm_b200.h")).read()
    m = re.search(r"#define\s+HMM_B200_NOALLOC_SCRATCH_BYTES\s+\((\d+)u\s*<<\s*(\d+)\)", h)
    assert m and (int(m.group(1)) << int(m.group(2))) == 64 << 20
    assert "no counterpart in the reference" in h.lower()
