"""Byte-compiles the reference's pure-Python layer (orion/: the compiler, orion.nn, the backend-agnostic wrappers under
orion/backend/python) where it lies under /root/reference into the archive oracle/_ref/orion_pyc.zip -- TEST INFRASTRUCTURE ONLY.

Same idea as `make -C oracle ref` for the C++ pieces: a build OUTPUT of the unmodified reference (CPython bytecode, no source
file is copied), git-ignored, travelling to the GPU box with the tree, so that tests/test_gpu_live.py can drive this backend
with the reference's own `orion.init_scheme / fit / compile / encode / encrypt / net.he() / net(ct)` flow.  Nothing outside
tests/ imports it.   python oracle/build_refpy.py
"""
import os
import py_compile
import sys

REF = "/root/reference"
HERE = os.path.dirname(os.path.abspath(__file__))
OUT = os.path.join(HERE, "_ref", "orion_pyc")
SKIP = ("backend/heongpu/adrianHEonGPU", "backend/lattigo/lattigo")       # vendored native trees: nothing Python to run


def build():
    src_root = os.path.join(REF, "orion")
    if not os.path.isdir(src_root):
        print("reference tree absent: keeping prebuilt oracle/_ref/orion_pyc (if any)")
        return False
    n = 0
    for d, _dirs, files in os.walk(src_root):
        rel = os.path.relpath(d, REF)
        if any(s in rel for s in SKIP):
            continue
        for f in files:
            if not f.endswith(".py"):
                continue
            dst = os.path.join(OUT, rel, f + "c")
            os.makedirs(os.path.dirname(dst), exist_ok=True)
            py_compile.compile(os.path.join(d, f), cfile=dst, dfile=os.path.join(rel, f), doraise=True, optimize=0)
            n += 1
    # directories the reference leaves as implicit namespace packages (orion/backend/heongpu) get an empty __init__:
    # zipimport does not resolve namespace packages
    empty = os.path.join(OUT, "_empty.py")
    open(empty, "w").close()
    for d, _dirs, files in os.walk(os.path.join(OUT, "orion")):
        if any(f.endswith(".pyc") for f in files) and "__init__.pyc" not in files:
            py_compile.compile(empty, cfile=os.path.join(d, "__init__.pyc"), dfile="__init__.py", doraise=True)
    os.remove(empty)
    # one archive (zipimport reads .pyc-only archives): loose *.pyc files are dropped from the snapshot that goes to the GPU box
    import shutil
    import zipfile
    with zipfile.ZipFile(OUT + ".zip", "w", zipfile.ZIP_DEFLATED) as z:
        for d, _dirs, files in os.walk(OUT):
            for f in files:
                full = os.path.join(d, f)
                z.write(full, os.path.relpath(full, OUT))
    shutil.rmtree(OUT)
    print("byte-compiled %d reference modules into %s.zip" % (n, OUT))
    return True


if __name__ == "__main__":
    sys.exit(0 if build() or os.path.exists(OUT + ".zip") else 1)
