"""CPU-only checks of the boundary: the C-ABI library builds/loads, exports every symbol the header
declares, and refuses to run without a GPU (no CPU fallback)."""
import os
import re

import pytest

import ragstore_b200 as rb

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared():
    src = open(os.path.join(ROOT, "include", "ragstore.h")).read()
    return sorted(set(re.findall(r"RAGSTORE_API\s+[\w\s\*]+?\b(ragstore_\w+)\s*\(", src)))


def test_header_declares_entry_points():
    names = _declared()
    for must in ("ragstore_create", "ragstore_add", "ragstore_search", "ragstore_count", "ragstore_destroy"):
        assert must in names


def test_library_exports_every_declared_symbol():
    L = rb.lib()
    for name in _declared():
        assert hasattr(L, name), f"{name} declared in include/ragstore.h but not exported"
    assert L.ragstore_version() >= 1000


def test_no_cpu_fallback_without_device():
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    with pytest.raises(rb.RagStoreError) as ei:
        rb.RagStore(8, 16)
    assert ei.value.code == 5  # RAGSTORE_ERR_NO_DEVICE


def test_product_does_not_import_oracle():
    pkg = os.path.join(ROOT, "llm-kubernetes-minikube-sharp4dev_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h", ".cpp")):
                txt = open(os.path.join(dirpath, f), errors="replace").read()
                assert not re.search(r"^\s*(from|import)\s+oracle\b", txt, re.M), f
                assert "liboracle" not in txt, f


def test_hit_filters_match_the_callers_of_the_reference():
    """N2: /rag/search low-score note (Program.cs:77-89) and /agent_rag citation rule (:125-128), host arithmetic only."""
    from ragstore_b200 import native
    scores = [0.71, 0.50, 0.43, 0.42, 0.10, -0.2]
    cited, best, low = native.hit_filters(scores)                 # max(0.35, 0.6 * 0.71 = 0.426)
    assert best == 0.71 and not low
    assert list(cited) == [True, True, True, False, False, False]
    cited, best, low = native.hit_filters([0.30, 0.19])            # best*0.6 = 0.18 < 0.35 -> nothing cited; not low (0.30 >= 0.20)
    assert list(cited) == [False, False] and best == 0.30 and not low
    cited, best, low = native.hit_filters([0.15, 0.12])
    assert low and best == 0.15
    cited, best, low = native.hit_filters([])
    assert list(cited) == [] and not low
    cited, best, low = native.hit_filters([0.9, 0.54, 0.5399999])  # boundary: >= 0.6*best
    assert list(cited) == [True, True, False]


def _build_c_smoke():
    import subprocess
    exe = os.path.join(ROOT, "tests", "c", "abi_smoke")
    so_dir = os.path.join(ROOT, "llm-kubernetes-minikube-sharp4dev_b200")
    subprocess.check_call(["gcc", "-std=c11", "-O1", "-o", exe, os.path.join(ROOT, "tests", "c", "abi_smoke.c"),
                           "-L" + so_dir, "-lragstore", "-Wl,-rpath," + so_dir])
    return exe


def test_c_abi_from_plain_c_without_gpu():
    """The header compiles as C and the library links from C; without a GPU it refuses cleanly (exit code 3)."""
    import subprocess
    import torch
    exe = _build_c_smoke()
    r = subprocess.run([exe], capture_output=True, text=True)
    if torch.cuda.is_available():
        assert r.returncode == 0, r.stderr
    else:
        assert r.returncode == 3, (r.returncode, r.stdout, r.stderr)


@pytest.mark.gpu
def test_c_abi_from_plain_c_on_gpu():
    import subprocess
    r = subprocess.run([_build_c_smoke()], capture_output=True, text=True)
    assert r.returncode == 0 and "abi_smoke ok" in r.stdout, (r.stdout, r.stderr)
