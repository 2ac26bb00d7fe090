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


def test_header_is_plain_c_and_a_c_host_links_and_calls(tmp_path):
    """The boundary is a C ABI: the header compiles as C11 with no C++ or torch types, and a C host binds the library
    through the DECLARED prototypes (a mismatch between a prototype and the exported symbol's use shows up here, not in
    ctypes).  The calls made need no GPU: version, error string, argument validation that fails before any CUDA call."""
    src = tmp_path / "host.c"
    src.write_text(r'''
#include <stdio.h>
#include <string.h>
#include "slider_b200.h"
int main(void) {
  if (slider_version() != 3) { printf("version %d\n", slider_version()); return 2; }
  slider_layer_geom g;
  memset(&g, 0, sizeof g);
  g.dtype = SLIDER_BF16; g.batch = 1; g.height = 1; g.width = 8; g.cin = 8; g.cout = 8;
  g.kh = 5; g.kw = 5; g.pad = 2; g.rank = 4;                       /* 5x5 taps: rejected by validation */
  int rc = slider_lora_forward(&g, (void*)16, (const void*)16, (const void*)16, 0, 0, 0, 0, 1.0f, 0, 0, 0);
  if (rc == SLIDER_OK) return 3;
  if (strlen(slider_last_error()) == 0) return 4;
  rc = slider_cross_attention((void*)16, (const void*)16, (const void*)16, (const void*)16, 1, 64, 4096, 8, 40, SLIDER_BF16, 0, 0);
  if (rc == SLIDER_OK) return 5;                                    /* 4096 keys: outside the resident-key kernel */
  printf("ok %s\n", slider_last_error());
  return 0;
}
''')
    exe = tmp_path / "host"
    lib = build.build()
    cmd = ["gcc", "-std=c11", "-Wall", "-Werror", "-pedantic", "-I", str(ROOT / "include"), str(src), "-o", str(exe),
           "-L", str(lib.parent), "-l:" + lib.name, "-Wl,-rpath," + str(lib.parent)]
    r = subprocess.run(cmd, capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    r = subprocess.run([str(exe)], capture_output=True, text=True)
    assert r.returncode == 0 and r.stdout.startswith("ok "), (r.returncode, r.stdout, r.stderr)


def test_missing_extension_fails_loudly(monkeypatch, tmp_path):
    """No fallback: with the shared library absent the first call through the binding raises and names the build command."""
    import pytest
    monkeypatch.setattr(_lib, "_lib", None)
    monkeypatch.setattr(_lib, "LIB_PATH", tmp_path / "libslider_b200.so")
    with pytest.raises(RuntimeError, match="no fallback"):
        _lib.lib()
    with pytest.raises(RuntimeError, match="no fallback"):
        from sliders_conceptmod_b200 import ops
        ops.lib().slider_version()
