"""The drop-in boundary without a GPU: libnsof.so (the product build, sm_100a) loads, exports every symbol
include/nsof.h declares, the ctypes table of the package covers exactly that set, and on a box without a CUDA device the
product path fails loudly instead of falling back to anything on the CPU."""
import ctypes as C
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared():
    src = open(os.path.join(ROOT, "include", "nsof.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"NSOF_API\s+[\w\s\*]+?\b(nsof_\w+)\s*\(", src)))


def test_library_exports_every_declared_symbol(nsof):
    names = _declared()
    assert len(names) >= 45 and "nsof_farneback_u8_host" in names and "nsof_flow_segment_batch" in names
    lib = C.CDLL(nsof.DEFAULT_LIB_PATH)  # loading needs libcudart, not a device
    missing = [n for n in names if not hasattr(lib, n)]
    assert not missing, missing
    assert sorted(nsof.EXPORTED_SYMBOLS) == names  # the package binds the whole header and nothing else


def test_no_cpu_fallback_without_a_device(nsof):
    import torch
    if torch.cuda.is_available():
        pytest.skip("a CUDA device is present")
    with pytest.raises(Exception) as e:
        nsof.Engine(0)
    assert "cuda" in str(e.value).lower() or "device" in str(e.value).lower()
    import numpy as np
    a = np.zeros((40, 40), np.uint8)
    with pytest.raises(Exception):
        nsof.calcOpticalFlowFarneback(a, a, None, **nsof.FARNEBACK_PARAMS)
