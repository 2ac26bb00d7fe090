"""The C-ABI library builds for sm_100a, loads without a GPU/driver and exports every symbol that
include/slider_b200.h declares.  No compute call is made here."""
import ctypes
import re
import subprocess
from pathlib import Path

from helpers import ROOT
from sliders_conceptmod_b200 import _lib, build


def _declared():
    text = (ROOT / "include" / "slider_b200.h").read_text()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(slider_[a-z0-9_]+)\s*\(", text)))


def test_library_builds_and_exports_header_symbols():
    path = build.build()
    assert path.exists()
    names = _declared()
    assert len(names) >= 15
    h = ctypes.CDLL(str(path))
    for n in names:
        assert hasattr(h, n), f"{n} declared in include/slider_b200.h but not exported"
    assert set(names) == set(_lib.EXPORTS), "ctypes signature table out of sync with the header"


def test_version_and_error_string_without_gpu():
    h = _lib.lib()
    assert h.slider_version() == 3
    assert isinstance(h.slider_last_error(), bytes)
    g = _lib.LayerGeom(1, 1, 1, 8, 8, 8, 5, 5, 2, 4)       # 5x5 kernel: rejected before any CUDA call
    rc = h.slider_lora_forward(ctypes.byref(g), 16, 16, 16, None, None, None, None, 1.0, None, 0, None)
    assert rc == 1 and b"1x1" in h.slider_last_error()


def test_sass_is_blackwell_native():
    for obj in ("lora_gemm_bn160.o", "lora_gemm_bn80.o"):       # the two tile widths of lora_gemm_kernel
        out = subprocess.run(["cuobjdump", "-sass", str(build.PKG_DIR / "build" / obj)], capture_output=True,
                             text=True).stdout
        for mnemonic in ("UTCHMMA", "UTMALDG", "UTMASTG", "LDTM"):
            assert mnemonic in out, f"{mnemonic} missing from the SASS of {obj}"
