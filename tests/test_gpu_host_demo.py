"""The plain-C host program (fluid-simulation_b200/host/fluid_host_demo.c): host code that stays in C and calls
only the extern "C" layer of include/fluid_cuda.h.  It is executed as a program and the field hashes it prints
are compared with the golden hashes of the unmodified reference (tests/golden/golden_hashes.json)."""
import os
import re
import subprocess

import pytest

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
DEMO = os.path.join(ROOT, "fluid-simulation_b200", "lib", "fluid_host_demo")


def run_demo(*args):
    assert os.path.exists(DEMO), "fluid_host_demo not built (make -C fluid-simulation_b200)"
    out = subprocess.run([DEMO] + [str(a) for a in args], capture_output=True, text=True, timeout=600)
    assert out.returncode == 0, out.stdout + out.stderr
    return dict(re.findall(r"^\s+(\w+)\s+fnv1a64=([0-9a-f]{16})", out.stdout, re.M)), out.stdout


def test_c_host_program_reproduces_the_reference_hashes(golden_hashes):
    for fr in golden_hashes:
        if fr["source"] != "stam" or fr["vorticity"] != 0.0 or fr["damp"] != 0.0 or fr["N"] * fr["N"] * fr["steps"] > 64 * 64 * 1000:
            continue  # the program drives source S1 without vorticity or damping
        got, text = run_demo(fr["N"], fr["steps"], repr(fr["visc"]), repr(fr["diff"]), 0)
        assert (got.get("dens"), got.get("u"), got.get("v")) == (fr["dens"], fr["u"], fr["v"]), (fr, text)


def test_c_host_program_red_black_and_ensemble_run():
    got, text = run_demo(256, 5, "1e-4", "1e-4", 1, 3)
    assert set(got) == {"u", "v", "dens"} and "red-black" in text
