"""include/b200labs.h against the built library: every function the header declares must be
exported by libb200labs.so (catches a renamed or dropped entry point), and every exported
b200_* symbol must be declared in the header (catches an undocumented one)."""
import re
import subprocess


def declared_functions(header_text: str):
    text = re.sub(r"/\*.*?\*/", " ", header_text, flags=re.S)          # drop comments
    text = re.sub(r"^\s*#.*$", " ", text, flags=re.M)                   # and preprocessor lines
    return set(re.findall(r"\b(b200_[a-z0-9_]+)\s*\(", text))


def test_header_and_library_agree(built, lib, repo_root):
    declared = declared_functions((repo_root / "include" / "b200labs.h").read_text())
    assert len(declared) > 40
    missing = sorted(name for name in declared if not hasattr(lib, name))
    assert not missing, f"declared in b200labs.h but not exported: {missing}"
    nm = subprocess.run(["nm", "-D", "--defined-only", str(repo_root / "cuda_mpi_openmp_b200/lib/libb200labs.so")],
                        capture_output=True, text=True, check=True).stdout
    exported = set(re.findall(r"\bT (b200_[a-z0-9_]+)$", nm, flags=re.M))
    undeclared = sorted(exported - declared)
    assert not undeclared, f"exported but not declared in b200labs.h: {undeclared}"
