"""The drop-in boundary without a GPU: libthetal_b200.so loads and exports every entry point that
include/thetal_b200.h declares, and e3sm_b200/lib.py lists the same set.  No compute call is made."""
import os
import re
import subprocess

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "thetal_b200.h")


def declared_symbols():
    text = open(HEADER).read()
    text = re.sub(r"/\*.*?\*/", " ", text, flags=re.S)          # comments
    text = re.sub(r"//[^\n]*", " ", text)
    text = re.sub(r"^\s*#[^\n]*", " ", text, flags=re.M)         # preprocessor lines
    names = []
    for decl in text.split(";"):
        if "(" not in decl or "typedef" in decl:
            continue
        m = re.search(r"([A-Za-z_][A-Za-z0-9_]*)\s*\(", decl)
        if m:
            names.append(m.group(1))
    return names


def library():
    from e3sm_b200 import lib
    if not os.path.exists(lib.LIB_PATH):
        subprocess.run(["make", "-C", os.path.dirname(lib.LIB_PATH)], check=True, capture_output=True)
    return lib.load()          # the product's own loader (it orders the NCCL copies, see lib._preload_bundled_nccl)


def test_header_declares_entry_points():
    names = declared_symbols()
    assert len(names) >= 40 and len(set(names)) == len(names)
    for must in ("prim_advance_exp_c", "caar_c", "dirk_c", "advance_hypervis_c", "init_simulation_params_c",
                 "thetal_b200_global_diagnostics", "thetal_b200_global_diagnostics_noalloc"):
        assert must in names, must


def test_library_exports_every_declared_symbol():
    L = library()
    missing = [n for n in declared_symbols() if not hasattr(L, n)]
    assert not missing, missing


def test_loader_lists_the_header():
    from e3sm_b200 import lib
    assert sorted(lib.SYMBOLS) == sorted(declared_symbols())


def test_torch_imports_after_the_library():
    """The library and PyTorch share the soname libnccl.so.2: loading the library first must not break `import torch`
    (a fresh interpreter, library first)."""
    import sys
    code = ("import sys; sys.path.insert(0, %r); from e3sm_b200 import lib; lib.load(); import torch; "
            "import torch.distributed; print('ok')" % ROOT)
    out = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, timeout=600)
    assert out.returncode == 0 and out.stdout.strip().endswith("ok"), out.stderr[-2000:]
