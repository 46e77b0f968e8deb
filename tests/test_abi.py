"""The C-ABI library builds, loads and exports every symbol include/abub_cuda.h declares
(no compute calls: there is no GPU on the CPU box)."""
import ctypes
import os
import re

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_library_exports_every_declared_symbol():
    from autobub3hs_b200 import build, capi
    lib_path = build.build()
    lib = ctypes.CDLL(lib_path)
    hdr = open(os.path.join(ROOT, "include", "abub_cuda.h")).read()
    hdr = re.sub(r"/\*.*?\*/", "", hdr, flags=re.S)
    declared = set(re.findall(r"\b(abub_[a-z0-9_]+)\s*\(", hdr))
    assert declared, "no declarations parsed"
    for name in sorted(declared):
        assert hasattr(lib, name), f"{name} declared in abub_cuda.h but not exported"
    assert declared == set(capi.EXPORTS)


def test_blob_struct_layout():
    from autobub3hs_b200 import capi
    assert ctypes.sizeof(capi.Blob) == 8 * 4 + 2 * 8 + 4 * 8


def test_host_library_exports_every_declared_symbol():
    """libabubhost.so (host classes + the harness entry points of include/abub_host.h) links and exports them."""
    import subprocess
    from autobub3hs_b200 import build, capi
    build.build()
    subprocess.check_call(["make", "-s", "-C", os.path.join(ROOT, "autobub3hs_b200", "host")])
    hl = capi.load_host()
    hdr = open(os.path.join(ROOT, "include", "abub_host.h")).read()
    hdr = re.sub(r"/\*.*?\*/", "", hdr, flags=re.S)
    declared = set(re.findall(r"\b(abubhost_[a-z0-9_]+)\s*\(", hdr))
    assert declared == set(capi.HOST_EXPORTS)
    for name in sorted(declared):
        assert hasattr(hl, name), f"{name} declared in abub_host.h but not exported"
