"""CPU suite, part 2: the C-ABI library loads without a GPU, exports every
symbol include/fluid_cuda.h declares, validates arguments, and -- with no CUDA
device -- fails loudly instead of falling back to the CPU."""
import ctypes
import os
import re
import subprocess

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "fluid_cuda.h")


def declared_symbols():
    text = open(HEADER).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(fluid_[a-z0-9_]+)\s*\(", text)))


def test_header_is_plain_c(tmp_path):
    src = tmp_path / "t.c"
    src.write_text('#include "fluid_cuda.h"\nint main(void){fluid_params p; fluid_params_default(&p); return (int)sizeof(p);}\n')
    subprocess.check_call(["gcc", "-std=c99", "-pedantic", "-Wall", "-Werror", "-I", os.path.join(ROOT, "include"),
                           "-c", str(src), "-o", str(tmp_path / "t.o")])


def test_library_exports_every_declared_symbol(fl):
    L = fl.lib()
    decl = declared_symbols()
    assert len(decl) >= 40
    missing = [s for s in decl if not hasattr(L, s)]
    assert not missing, missing
    assert sorted(fl.EXPORTS) == decl, set(fl.EXPORTS) ^ set(decl)


def test_library_does_not_link_the_oracle(fl):
    out = subprocess.check_output(["ldd", fl.LIB_PATH]).decode()
    assert "oracle" not in out and "fluid_ref" not in out
    syms = subprocess.check_output(["nm", "-D", fl.LIB_PATH]).decode()
    assert "oracle_" not in syms and "ref_update" not in syms


def test_params_struct_layout(fl):
    p = fl.Params()
    fl.lib().fluid_params_default(ctypes.byref(p))
    assert (p.N, p.batch, p.gs_iters, p.gs_order, p.advect_mode, p.device) == (64, 1, 20, 0, 0, -1)
    assert ctypes.sizeof(fl.Params) == 16 * 4


def _has_gpu():
    try:
        import torch
        return torch.cuda.is_available()
    except Exception:
        return False


def test_argument_validation(fl):
    L = fl.lib()
    assert L.fluid_set_gs_order(7) == 1 and b"gs order" in L.fluid_last_error()
    assert L.fluid_set_gs_iters(0) == 1
    assert L.fluid_set_advect_mode(-1) == 1
    p = fl.Params()
    L.fluid_params_default(ctypes.byref(p))
    h = ctypes.c_void_p()
    p.N = 0
    assert L.fluid_ctx_create(ctypes.byref(p), ctypes.byref(h)) == 1
    p.N, p.batch = 8, 0
    assert L.fluid_ctx_create(ctypes.byref(p), ctypes.byref(h)) == 1
    assert L.fluid_ctx_step(None, 0.1, 1) == 1 and b"null context" in L.fluid_last_error()


@pytest.mark.skipif(_has_gpu(), reason="a CUDA device is present")
def test_no_device_means_loud_failure_not_cpu_fallback(fl):
    x = np.ones(16, np.float32)
    with pytest.raises(fl.FluidError) as e:
        fl.add_source(2, x, x.copy(), 0.1)
    assert e.value.code == 2 and "no CPU fallback" in str(e.value)
    assert np.all(x == 1)  # untouched
    with pytest.raises(fl.FluidError):
        fl.FluidSystem().Setup(8)
