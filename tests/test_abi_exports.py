"""CPU-only checks of the drop-in boundary: the CUDA library loads and exports every function include/hsq_b200.h
declares; the ctypes struct mirrors have the sizes the header implies; the product package never imports the oracle."""
import ctypes as C
import os
import re

from bioinformatics_b200 import _abi as A
from bioinformatics_b200 import _lib, build

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared():
    text = open(os.path.join(ROOT, "include", "hsq_b200.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(hsq_[a-z0-9_]+)\s*\(", text)))


def test_library_exports_every_declared_function():
    build.build_native()
    lib = C.CDLL(_lib.SO_PATH)
    names = _declared()
    assert len(names) >= 20
    for name in names:
        assert hasattr(lib, name), "libhsq_b200.so does not export %s" % name
    assert set(_lib.EXPORTS) == set(names)


def test_every_export_has_a_java_handle():
    """integration/java/HsqB200.java (the reference-side binding, Panama FFM) binds every function the header declares."""
    java = open(os.path.join(ROOT, "integration", "java", "HsqB200.java")).read()
    bound = set(re.findall(r'fn\("(hsq_[a-z0-9_]+)"', java))
    missing = [n for n in _declared() if n not in bound]
    assert not missing, "no Java downcall handle for %s" % missing
    # the operator seam and the two batch seams exist as source
    for f in ("GpuNeedlemanWunschGlobalAlignment.java", "HsqBatcher.java", "FilterByUidGpu.java", "ExtendReadsToPrimerGpu.java"):
        assert os.path.exists(os.path.join(ROOT, "integration", "java", f)), f


def test_struct_sizes():
    assert C.sizeof(A.Scorer) == 16
    assert C.sizeof(A.Seqs) == 24
    assert C.sizeof(A.PairBatch) == 8 + 8 + 8 + 24 + 24 + 8 + 8
    assert C.sizeof(A.ProbeStats) == 32


def test_no_device_is_an_error_not_a_fallback():
    """Without a GPU hsq_ctx_create must fail loudly (this test only asserts that on machines without one)."""
    import torch
    if torch.cuda.is_available():
        return
    lib = _lib.load()
    h = C.c_void_p()
    st = lib.hsq_ctx_create(C.c_int32(0), C.byref(h))
    assert st == A.HSQ_ERR_NO_DEVICE and not h.value
    assert b"no CPU fallback" in lib.hsq_last_error(None)


def test_product_package_does_not_touch_the_oracle():
    """No file of the product package imports, includes or loads anything under oracle/ (prose mentions are fine)."""
    pkg = os.path.join(ROOT, "bioinformatics_b200")
    pat = re.compile(r"import\s+oracle|from\s+oracle|#include\s+\"[^\"]*oracle|libhsq_oracle|oracle/")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".inl", ".h")):
                src = open(os.path.join(dirpath, f), errors="replace").read()
                assert not pat.search(src), "%s reaches into the oracle" % f
