"""The drop-in boundary: libptb200.so loads, exports every entry point include/ptb200.h declares, and — on a
box without a GPU — refuses to work instead of falling back to a CPU path."""
import ctypes
import os
import re

import pytest

from opencl_path_tracer_b200 import capi
from tests import util

HEADER = os.path.join(util.ROOT, "include", "ptb200.h")


def declared_symbols():
    text = open(HEADER).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(ptb_[a-z0-9_]+)\s*\(", text)))


def test_library_is_built():
    assert os.path.exists(capi.lib_path()), "run __graft_entry__.build() (make -C opencl_path_tracer_b200/csrc)"


def test_every_declared_symbol_is_exported_and_bound():
    lib = ctypes.CDLL(capi.lib_path())
    names = declared_symbols()
    assert len(names) >= 40
    for n in names:
        assert hasattr(lib, n), f"{n} declared in include/ptb200.h but not exported"
    assert sorted(capi.SYMBOLS) == names, "capi.SYMBOLS and the header disagree"
    capi.load_library()     # prototypes for all of them


def test_kernel_names_resolve_without_a_device():
    lib = capi.load_library()
    assert lib.ptb_kernel_id(b"gen_ray") == 0 and lib.ptb_kernel_id(b"trace_ray") == 1 and lib.ptb_kernel_id(b"filt_im") == 2
    assert lib.ptb_kernel_id(b"nope") < 0
    assert lib.ptb_backend_name() == b"cuda-sm100a"


def test_no_cpu_fallback_without_gpu():
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    with pytest.raises(capi.PtbError, match="no CUDA device"):
        capi.Ptb(0)


def test_product_does_not_import_the_oracle():
    # the oracle is test infrastructure: nothing under the package may reference it
    pkg = os.path.join(util.ROOT, "opencl_path_tracer_b200")
    for root, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h", ".hpp", ".cpp")):
                src = open(os.path.join(root, f), errors="ignore").read()
                for needle in ("pt_oracle", "libpt_oracle", "ref_cpu", "import oracle", "from oracle", "ref_backend"):
                    assert needle not in src, (needle, os.path.join(root, f))
