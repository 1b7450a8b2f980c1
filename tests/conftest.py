"""Shared fixtures.

`-m "not gpu"` (CPU, this container and the driver's CPU tier): oracle vs golden
vectors, host front end vs reference dumps, C-ABI symbol checks, device-function
logic stepped on the CPU (tests/hostsim).
`-m gpu` (B200): the parity tests proper, through the C ABI of libpbrtgpu.so.
"""
import ctypes as C
import importlib
import os
import subprocess
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))

REF_DIR = os.path.join(ROOT, "oracle", "_ref")
REF_SCENES = os.path.join(REF_DIR, "scenes")
GOLDEN = os.path.join(ROOT, "tests", "golden")
TEST_SCENES = os.path.join(ROOT, "tests", "scenes")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session")
def pkg():
    mod = importlib.import_module("pbrt-v3-light-portals_b200")
    if not os.path.exists(os.path.join(ROOT, "pbrt-v3-light-portals_b200", "lib", "libpbrthost.so")):
        mod.build()
    return mod


@pytest.fixture(scope="session")
def oracle_lib():
    """The CPU restatement (oracle/port), built on demand."""
    path = os.path.join(ROOT, "oracle", "_build", "liboracle_port.so")
    srcs = [os.path.join(ROOT, "oracle", "port", f) for f in os.listdir(os.path.join(ROOT, "oracle", "port"))]
    if not os.path.exists(path) or any(os.path.getmtime(s) > os.path.getmtime(path) for s in srcs):
        subprocess.check_call(["make", "-s", "-C", os.path.join(ROOT, "oracle"), "port"])
    return C.CDLL(path)


@pytest.fixture(scope="session")
def hostsim_lib():
    """Device headers compiled for the CPU (tests/hostsim) -- a debugging aid."""
    out = os.path.join(ROOT, "tests", "_build", "libhostsim.so")
    src = os.path.join(ROOT, "tests", "hostsim", "hostsim.cpp")
    dev = os.path.join(ROOT, "pbrt-v3-light-portals_b200", "csrc")
    deps = [src] + [os.path.join(dp, f) for dp, _, fs in os.walk(dev) for f in fs if f.endswith(".h")]
    if not os.path.exists(out) or any(os.path.getmtime(d) > os.path.getmtime(out) for d in deps):
        os.makedirs(os.path.dirname(out), exist_ok=True)
        subprocess.check_call(["g++", "-std=c++17", "-O2", "-ffp-contract=off", "-fPIC", "-shared", "-pthread", src, "-o", out])
    lib = C.CDLL(out)
    lib.hostsim_create.restype = C.c_void_p
    lib.hostsim_create.argtypes = [C.c_void_p]
    lib.hostsim_destroy.argtypes = [C.c_void_p]
    return lib


@pytest.fixture(scope="session")
def hostsim_perturbed_lib():
    """tests/hostsim built with -DPG_PERTURB_LIBM: sinf / cosf / acosf / atan2f one ulp off in half of the calls."""
    out = os.path.join(ROOT, "tests", "_build", "libhostsim_perturbed.so")
    src = os.path.join(ROOT, "tests", "hostsim", "hostsim.cpp")
    dev = os.path.join(ROOT, "pbrt-v3-light-portals_b200", "csrc")
    deps = [src] + [os.path.join(dp, f) for dp, _, fs in os.walk(dev) for f in fs if f.endswith(".h")]
    if not os.path.exists(out) or any(os.path.getmtime(d) > os.path.getmtime(out) for d in deps):
        os.makedirs(os.path.dirname(out), exist_ok=True)
        subprocess.check_call(["g++", "-std=c++17", "-O2", "-ffp-contract=off", "-fPIC", "-shared", "-pthread", "-DPG_PERTURB_LIBM", src, "-o", out])
    lib = C.CDLL(out)
    lib.hostsim_create.restype = C.c_void_p
    lib.hostsim_create.argtypes = [C.c_void_p]
    lib.hostsim_destroy.argtypes = [C.c_void_p]
    return lib


@pytest.fixture(scope="session")
def ref_harness():
    """oracle/_ref/ref_harness (the reference's own code); None when not built."""
    path = os.path.join(REF_DIR, "ref_harness")
    return path if os.path.exists(path) else None


@pytest.fixture(scope="session")
def gpu(pkg):
    ctx = pkg.GpuContext(0)
    yield ctx
    ctx.close()


def scene_path(name):
    """Test scenes live in tests/scenes; reference scenes are staged in oracle/_ref/scenes."""
    p = os.path.join(TEST_SCENES, name)
    if os.path.exists(p):
        return p
    p = os.path.join(REF_SCENES, name)
    if os.path.exists(p):
        return p
    pytest.skip("scene %s not staged (run `make -C oracle scenes` where /root/reference exists)" % name)
