"""CPU suite: the C-ABI shared library loads and exports every symbol include/igpu_rollup.h declares
(no compute without a GPU), and refuses loudly to run without a CUDA device (no CPU fallback)."""
import ctypes
import os
import re

import pytest

from headlamp_intel_gpu_plugin_b200 import native

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "igpu_rollup.h")


def declared_functions():
    src = open(HEADER).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(igpu_[a-z0-9_]+)\s*\(", src)))


def test_header_and_binding_list_agree():
    assert declared_functions() == sorted(native.EXPORTS)


def test_library_exports_every_declared_symbol():
    assert os.path.exists(native.LIB_PATH), "libigpu_rollup.so not built: run __graft_entry__.build()"
    lib = ctypes.CDLL(native.LIB_PATH)
    for name in declared_functions():
        assert hasattr(lib, name), f"{name} declared in include/igpu_rollup.h but not exported"


def test_version_and_strerror():
    lib = native.load_library()
    assert lib.igpu_version() == (0 << 16) | (6 << 8) | 0  # include/igpu_rollup.h IGPU_VERSION_*
    assert lib.igpu_strerror(0) == b"ok"
    assert b"no CPU fallback" in lib.igpu_strerror(-5)


def test_struct_layouts_match_header():
    # sizes the C side static-asserts implicitly: 16-byte entries, 24-byte node stats
    from headlamp_intel_gpu_plugin_b200 import columns as K

    assert K.ENTRY_DTYPE.itemsize == 16 and K.ENTRY_DTYPE.fields["value"][1] == 8
    assert K.NODE_STATS_DTYPE.itemsize == 24
    assert ctypes.sizeof(native.FleetColsC) == 24 + 10 * 8
    assert ctypes.sizeof(native.FleetOutC) == 20 * 8
    assert K.CONTAINER_ROW_DTYPE.itemsize == 32 and K.ENTRY_DTYPE.fields["container"][1] == 7
    assert ctypes.sizeof(native.TimingC) == 16 + 24 + 48


def test_no_device_means_error_not_fallback():
    import torch

    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    lib = native.load_library()
    h = ctypes.c_void_p()
    assert lib.igpu_ctx_create(0, None, ctypes.byref(h)) == -5  # IGPU_ERR_NODEVICE
    with pytest.raises(native.IgpuError):
        native.Context(0)
    from headlamp_intel_gpu_plugin_b200 import k8s

    native._default = None
    with pytest.raises(native.IgpuError):
        k8s.getNodeGpuCount({"metadata": {"name": "n"}, "status": {"capacity": {"gpu.intel.com/i915": "4"}}})


def test_missing_library_fails_loudly(tmp_path):
    with pytest.raises(native.NativeUnavailable):
        native.load_library(str(tmp_path / "nope.so"))


def test_product_never_imports_the_oracle():
    pkg = os.path.join(ROOT, "headlamp_intel_gpu_plugin_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".cpp", ".h")):
                text = open(os.path.join(dirpath, f), errors="ignore").read()
                assert "import oracle" not in text and "from oracle" not in text and "rollup_oracle" not in text, f
