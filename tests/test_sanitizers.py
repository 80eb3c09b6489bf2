"""Host-side sanitizer builds: the CPU oracle under AddressSanitizer + UndefinedBehaviorSanitizer, the
coalescing queue of ragstore_search_coalesced under ThreadSanitizer (GPU search stubbed), and the plain-C ABI smoke test
built with ASan + UBSan (on a box without a GPU it exercises the refusal path; `-m gpu` runs the real thing).  CPU only."""
import os
import shutil
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
BUILD = os.path.join(ROOT, "tests", "c", "build")


def _cc(args, out):
    os.makedirs(BUILD, exist_ok=True)
    exe = os.path.join(BUILD, out)
    r = subprocess.run(args + ["-o", exe], capture_output=True, text=True)
    if r.returncode != 0 and ("cannot find" in r.stderr or "unrecognized" in r.stderr):
        pytest.skip("this toolchain has no such sanitizer runtime: " + r.stderr.strip().splitlines()[-1])
    assert r.returncode == 0, r.stderr
    return exe


def _run(exe, env_extra=None, ok_codes=(0,)):
    env = dict(os.environ, ASAN_OPTIONS="detect_leaks=1:abort_on_error=0", UBSAN_OPTIONS="halt_on_error=1:print_stacktrace=1")
    env.update(env_extra or {})
    r = subprocess.run([exe], capture_output=True, text=True, env=env, timeout=300)
    assert r.returncode in ok_codes, (r.returncode, r.stdout[-2000:], r.stderr[-4000:])
    assert "ERROR: AddressSanitizer" not in r.stderr and "runtime error:" not in r.stderr and "WARNING: ThreadSanitizer" not in r.stderr, r.stderr[-4000:]
    return r


def test_oracle_under_asan_and_ubsan():
    exe = _cc(["gcc", "-std=c11", "-O1", "-g", "-fno-omit-frame-pointer", "-ffp-contract=off", "-fsanitize=address,undefined",
               os.path.join(ROOT, "tests", "c", "oracle_selftest.c"), os.path.join(ROOT, "oracle", "ragindex_oracle.c"), "-lm", "-lpthread"],
              "oracle_selftest_asan")
    assert "oracle_selftest ok" in _run(exe).stdout


def test_coalescing_queue_under_tsan():
    exe = _cc(["g++", "-std=c++17", "-O1", "-g", "-fsanitize=thread", os.path.join(ROOT, "tests", "c", "coalesce_stress.cpp"), "-lpthread"],
              "coalesce_stress_tsan")
    # TSan needs ASLR it can live with; some sandboxes refuse (exit code 66 "unexpected memory mapping") -- that is not a finding
    env = dict(os.environ, TSAN_OPTIONS="halt_on_error=1")
    r = subprocess.run([exe], capture_output=True, text=True, env=env, timeout=600)
    if r.returncode != 0 and "unexpected memory mapping" in r.stderr:
        if shutil.which("setarch"):
            r = subprocess.run(["setarch", os.uname().machine, "-R", exe], capture_output=True, text=True, env=env, timeout=600)
        if r.returncode != 0 and "unexpected memory mapping" in r.stderr:
            pytest.skip("ThreadSanitizer cannot map its shadow memory in this sandbox")
    assert r.returncode == 0 and "coalesce_stress ok" in r.stdout, (r.returncode, r.stdout[-1000:], r.stderr[-4000:])
    assert "WARNING: ThreadSanitizer" not in r.stderr


def test_coalescing_queue_plain():
    """The same stress without a sanitizer (always runnable): answers, batch kinds, nobody kept past its own answer."""
    exe = _cc(["g++", "-std=c++17", "-O2", os.path.join(ROOT, "tests", "c", "coalesce_stress.cpp"), "-lpthread"], "coalesce_stress")
    assert "coalesce_stress ok" in _run(exe).stdout


def test_dealing_arithmetic_under_ubsan():
    """csrc/dealing.h: how a multi-device handle deals rows to shards (ids = insertion order, prefixes stay prefixes)."""
    exe = _cc(["g++", "-std=c++17", "-O1", "-g", "-fsanitize=address,undefined", os.path.join(ROOT, "tests", "c", "dealing_selftest.cpp")],
              "dealing_selftest_ubsan")
    assert "dealing_selftest ok" in _run(exe).stdout


def _abi_smoke_asan():
    so_dir = os.path.join(ROOT, "llm-kubernetes-minikube-sharp4dev_b200")
    return _cc(["gcc", "-std=c11", "-O1", "-g", "-fsanitize=address,undefined", os.path.join(ROOT, "tests", "c", "abi_smoke.c"),
                "-L" + so_dir, "-lragstore", "-Wl,-rpath," + so_dir], "abi_smoke_asan")


def test_abi_smoke_under_asan_without_gpu():
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present: covered by the -m gpu variant")
    # protect_shadow_gap=0: the CUDA runtime inside libragstore.so reserves address space ASan would otherwise claim
    _run(_abi_smoke_asan(), {"ASAN_OPTIONS": "protect_shadow_gap=0:detect_leaks=0"}, ok_codes=(3,))


@pytest.mark.gpu
def test_abi_smoke_under_asan_on_gpu():
    r = _run(_abi_smoke_asan(), {"ASAN_OPTIONS": "protect_shadow_gap=0:detect_leaks=0"})
    assert "abi_smoke ok" in r.stdout
