"""CPU: the C-ABI library loads, exports every symbol include/aof.h declares, and
refuses to work (loudly, no CPU fallback) when no CUDA device is present."""
import ctypes as C
import os
import re
import subprocess

import __graft_entry__ as entry

ROOT = entry.ROOT


def declared_functions():
    names = []
    for hdr in sorted(os.listdir(os.path.join(ROOT, "include"))):
        text = open(os.path.join(ROOT, "include", hdr)).read()
        text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
        names += re.findall(r"\b(aof_[a-z0-9_]+)\s*\(", text)
    return sorted(set(names))


def test_every_declared_symbol_is_exported(pkg):
    L = pkg.lib()
    names = declared_functions()
    assert len(names) >= 14
    for n in names:
        assert hasattr(L, n), n


def test_product_does_not_link_oracle(pkg):
    out = subprocess.run(["ldd", pkg.capi.LIB_PATH], capture_output=True, text=True).stdout
    assert "oracle" not in out and "aof_ref" not in out
    syms = subprocess.run(["nm", "-D", pkg.capi.LIB_PATH], capture_output=True, text=True).stdout
    assert "orc_" not in syms and "ref_" not in syms
    pat = re.compile(r"^\s*(import\s+oracle|from\s+oracle|#\s*include\s*[<\"].*oracle)", re.M)
    for root, _, files in os.walk(entry.PKG_DIR):
        for f in files:
            if f.endswith((".cu", ".cuh", ".h", ".hpp", ".cpp", ".py")):
                assert not pat.search(open(os.path.join(root, f)).read()), f


def test_no_cpu_fallback_without_gpu(pkg):
    L = pkg.lib()
    if L.aof_device_count() > 0:
        return  # on the GPU box this property is not observable
    p = pkg.capi.Px4Params()
    L.aof_px4_default_params(C.byref(p))
    assert (p.output_rate, p.image_width, p.image_height, p.search_size) == (15, 64, 64, 6)
    p.focal_length_x = p.focal_length_y = 200.0
    h = C.c_void_p()
    rc = L.aof_px4_create(C.byref(p), 1, 0, C.byref(h))
    assert rc in (-102, -103) and not h.value
    assert L.aof_last_error()
    try:
        pkg.Px4Flow(200.0, 200.0)
    except pkg.AofError:
        pass
    else:
        raise AssertionError("Px4Flow must raise without a CUDA device")


def test_bad_geometry_rejected(pkg):
    L = pkg.lib()
    h = C.c_void_p()
    for s in (0, -3, 128):                                           # outside the reference's int8_t displacement loop
        p = pkg.capi.Px4Params(200.0, 200.0, 15, 64, 64, s, 30, 3000)
        assert L.aof_px4_create(C.byref(p), 1, 0, C.byref(h)) == -101
    p = pkg.capi.Px4Params(200.0, 200.0, 15, 16, 16, 4, 30, 3000)   # too small for the grid
    assert L.aof_px4_create(C.byref(p), 1, 0, C.byref(h)) == -101
    assert L.aof_px4_create(None, 1, 0, C.byref(h)) == -100
