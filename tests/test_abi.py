"""C-ABI surface checks that need no GPU: the product library loads, exports every symbol that
include/lora_phy_cuda.h declares, the ctypes table mirrors the header one to one, and without a
CUDA device construction fails loudly (-ENODEV) instead of falling back to anything."""
import ctypes as C
import errno
import os
import re

import pytest


def header_symbols(lp):
    txt = open(lp.HEADER).read()
    txt = re.sub(r"/\*.*?\*/", "", txt, flags=re.S)
    return sorted(set(re.findall(r"\b(lora_phy_cuda_[a-z_0-9]+)\s*\(", txt)))


@pytest.fixture(scope="module")
def lp():
    import lora_phy_b200
    if not os.path.exists(lora_phy_b200.LIB_PATH):
        lora_phy_b200.build()
    return lora_phy_b200


def test_library_exports_every_declared_symbol(lp):
    lib = C.CDLL(lp.LIB_PATH)
    names = header_symbols(lp)
    assert len(names) >= 20
    for n in names:
        assert hasattr(lib, n), f"{n} declared in include/lora_phy_cuda.h but not exported"


def test_ctypes_table_mirrors_header(lp):
    assert sorted(lp.SIGNATURES) == header_symbols(lp)
    lib = lp.load_library()
    assert lib.lora_phy_cuda_abi_version() == 1


def test_cfg_struct_layout(lp):
    # lora_phy_cuda_cfg is 8 x 32-bit fields (include/lora_phy_cuda.h)
    assert C.sizeof(lp.Cfg) == 32
    assert [f[0] for f in lp.Cfg._fields_] == ["sf", "bw_hz", "osr", "window", "flags", "sync_word", "amplitude", "reserved"]


def test_no_fallback_without_gpu(lp):
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    lib = lp.load_library()
    h = C.c_void_p()
    rc = lib.lora_phy_cuda_create(C.byref(h), 0)
    assert rc == -errno.ENODEV and not h.value
    with pytest.raises(lp.LoraPhyError):
        lp.LoraPhyCuda(0)
    # NULL-context calls report -EINVAL rather than crashing
    assert lib.lora_phy_cuda_synchronize(None) == -errno.EINVAL
    assert lib.lora_phy_cuda_last_error(None) == b"no context"


def test_product_never_references_the_oracle(lp):
    """The oracle is test infrastructure: no product source may mention it, and the built library
    must not link it."""
    pkg = os.path.dirname(lp.LIB_PATH)
    for root, _, files in os.walk(pkg):
        if "build" in root.split(os.sep):
            continue
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h", ".hpp", ".cpp")):
                src = open(os.path.join(root, f), errors="ignore").read()
                assert "liblora_oracle" not in src and "oracle_" not in src and "cpu_libs" not in src, f
    import subprocess
    out = subprocess.run(["ldd", lp.LIB_PATH], capture_output=True, text=True).stdout
    assert "oracle" not in out and "ref_lora" not in out
