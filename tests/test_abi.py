"""CPU: the C-ABI library loads and exports every symbol include/alegpu.h declares; host-side
argument validation (which needs no GPU) behaves like the reference's exceptions."""
import ctypes as C
import os
import re

import pytest

import alegpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_symbols():
    src = open(os.path.join(ROOT, "include", "alegpu.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(alegpu_[a-z0-9_]+)\s*\(", src)))


def test_library_exports_every_declared_symbol():
    lib = alegpu.lib()
    names = declared_symbols()
    assert len(names) >= 40
    missing = [n for n in names if not hasattr(lib, n)]
    assert not missing, missing
    assert b"sm_100a" in lib.alegpu_version()


def test_sass_is_sm100a_only():
    import subprocess
    out = subprocess.run(["cuobjdump", "-lelf", alegpu.SO_PATH], capture_output=True, text=True).stdout
    archs = set(re.findall(r"sm_\d+a?", out))
    assert archs == {"sm_100a"}, archs


def test_host_side_validation_matches_reference_exceptions():
    lib = alegpu.lib()
    # Buffer::get on a non-full buffer -> std::runtime_error (buffer.cc:64-65); checked before any CUDA call
    rc = lib.alegpu_buffer_get(None, None, None, None, None, None, C.c_int64(1), C.c_int64(1),
                               C.c_float(.99), C.c_float(.95), C.c_int64(3), None, None, None, 0,
                               None, None)
    assert rc == 2 and b"not full" in lib.alegpu_last_error()
    with pytest.raises(ValueError):
        alegpu._check(lib.alegpu_gae(None, None, None, None, None, None, None, C.c_int64(2),
                                     C.c_int64(2), C.c_float(.99), C.c_float(.95), None))
    with pytest.raises(ValueError):
        alegpu._check(lib.alegpu_loss(None, None, None, None, None, None, None, C.c_int64(4),
                                      C.c_int64(99), C.c_float(.1), C.c_float(.5), C.c_float(.01),
                                      None, None, None, None, None, None, None, None, None))


def test_no_cpu_fallback_without_device():
    if alegpu.device_count() > 0:
        pytest.skip("a GPU is present")
    with pytest.raises(alegpu.CudaError):
        alegpu.Learner(32, 4, 16)


def _build_shim_test(tmp_path):
    import subprocess
    exe = str(tmp_path / "shim_test")
    pkg = os.path.dirname(alegpu.SO_PATH)
    cmd = ["g++", "-std=c++17", "-O1", "-Wall", "-Werror", "-I" + os.path.join(ROOT, "include"),
           os.path.join(ROOT, "tests", "cpp", "shim_test.cc"), "-o", exe, "-L" + pkg, "-lalegpu",
           "-Wl,-rpath," + pkg]
    r = subprocess.run(cmd, capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    return exe


def test_cpp_shim_mirrors_reference_api(tmp_path):
    """include/ai/alegpu_shim.hpp (ai::gae::gae, ai::buffer::Buffer, ai::ppo::losses::compute,
    ai::ppo::train::train with the reference's names / argument order) compiles against the C ABI
    and throws the reference's exception types for the reference's conditions."""
    import subprocess
    exe = _build_shim_test(tmp_path)
    r = subprocess.run([exe], capture_output=True, text=True)
    assert r.returncode == 0 and "shim ok" in r.stdout, r.stdout + r.stderr


@pytest.mark.gpu
def test_cpp_shim_gae_known_answer_on_gpu(tmp_path):
    """test/ai/gae-test.cc:6-38 SimpleNoTerminals through the C++ shim on the device."""
    import subprocess
    exe = _build_shim_test(tmp_path)
    r = subprocess.run([exe, "gpu"], capture_output=True, text=True)
    assert r.returncode == 0 and "gpu gae: 2.81" in r.stdout, r.stdout + r.stderr
