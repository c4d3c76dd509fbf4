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


# Register budgets of the hot kernels = the CTAs per SM their measured times depend on (65 536 registers per SM).  A
# branch added to k_fb_um_first once cost it 6 registers, the fourth CTA per SM and 11 % of its time without any test
# noticing (profiles/r1v12_variant_ab.txt); this reads the budgets back from the product build.
REG_BUDGET = {
    r"k_fb_um_first_gILi0E": 64,                      # 4 CTAs x 256 threads
    r"k_fb_pyr_polyexp_tILi0ELi5ELi64ELi32E": 64,     # __launch_bounds__(256, 4)
    r"k_fb_pyr_polyexp_tILi1ELi5ELi64ELi32E": 64,
    r"k_fb_blur_solve_tILi7ELi64ELi32ELi256ELi3E": 80,  # 3 CTAs x 256 threads (shared memory allows no more)
    r"k_fb_blur_solve_gILi0E": 72,                    # box window of the generic kernel
}


def test_hot_kernels_keep_their_register_budget(nsof):
    import shutil
    import subprocess
    tool = shutil.which("cuobjdump") or "/usr/local/cuda/bin/cuobjdump"
    if not os.path.exists(tool):
        pytest.skip("cuobjdump not found")
    out = subprocess.run([tool, "--dump-resource-usage", nsof.DEFAULT_LIB_PATH], capture_output=True, text=True).stdout
    usage = dict(re.findall(r"Function (\S+):\s*\n\s*REG:(\d+)", out))
    for pat, budget in REG_BUDGET.items():
        hits = {k: int(v) for k, v in usage.items() if pat in k}
        assert hits, f"kernel {pat} not found in libnsof.so"
        for name, regs in hits.items():
            assert regs <= budget, f"{name}: {regs} registers > budget {budget} (occupancy regression)"
