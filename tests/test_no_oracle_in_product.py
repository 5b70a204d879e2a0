"""The oracle is test infrastructure: nothing under the product package may import it, call a CPU
fallback, or read the reference tree; bench.py may use it only in its CPU-baseline legs."""
import os
import re

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = os.path.join(ROOT, "vllm_from_scratch_b200")


def _sources(folder, exts):
    for dirpath, _, files in os.walk(folder):
        if "build" in dirpath.split(os.sep):
            continue
        for f in files:
            if f.endswith(exts):
                yield os.path.join(dirpath, f)


def test_package_never_imports_the_oracle_or_the_reference():
    bad = []
    for path in _sources(PKG, (".py", ".cu", ".cuh", ".h")):
        text = open(path, encoding="utf-8").read()
        if re.search(r"^\s*(from|import)\s+oracle\b", text, re.M) or "paged_attention_oracle" in text:
            bad.append((path, "imports the oracle"))
        if re.search(r"sys\.path[^\n]*reference|open\([^\n]*/root/reference", text):
            bad.append((path, "reads the reference tree"))
    assert not bad, bad


def test_bench_uses_the_oracle_only_in_cpu_legs():
    text = open(os.path.join(ROOT, "bench.py"), encoding="utf-8").read()
    # every oracle import sits inside one of the CPU-leg functions
    for m in re.finditer(r"from oracle import", text):
        head = text[:m.start()]
        func = re.findall(r"^def (\w+)\(", head, re.M)[-1]
        assert func in ("cpu_baseline", "_ref_worker", "run_reference"), func


def test_library_loader_has_no_fallback():
    text = open(os.path.join(PKG, "_lib.py"), encoding="utf-8").read()
    assert "raise RuntimeError" in text and "fallback" in text.lower()
    host = open(os.path.join(PKG, "paged_attention.py"), encoding="utf-8").read()
    assert "no CPU fallback" in host
