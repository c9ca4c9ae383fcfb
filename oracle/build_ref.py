"""Recipe that builds the checkers under ``oracle/``.  TEST INFRASTRUCTURE ONLY.

Two artefacts:

1. ``oracle/_ref/balltree*.so`` -- the UNMODIFIED reference ball-tree extension
   (``/root/reference/models/components/balltree/{balltree.pyx,balltree.h}``),
   cythonized and compiled from the sources where they lie.  Nothing is copied
   into the repository: the generated C++ goes to ``oracle/_ref/`` (git-ignored)
   and is deleted after the compile.  Only possible where ``/root/reference``
   exists (this container); the GPU box uses the prebuilt ``.so``.
   Mirrors the reference's own ``setup.py:9-30`` flags (-O3 -std=c++11 -fopenmp).

2. ``oracle/libballtree_oracle.so`` -- our plain-C restatement
   (``oracle/balltree_oracle.c``), compiled with gcc.

Nothing in the product package (``haet_b200/``) imports this module.
"""
from __future__ import annotations

import os
import shutil
import subprocess
import sys
import sysconfig

HERE = os.path.dirname(os.path.abspath(__file__))
REF_DIR = "/root/reference/models/components/balltree"
OUT_REF = os.path.join(HERE, "_ref")
CC = "/usr/bin/gcc" if os.path.exists("/usr/bin/gcc") else "gcc"
CXX = "/usr/bin/g++" if os.path.exists("/usr/bin/g++") else "g++"


def _newer(src: str, dst: str) -> bool:
    return (not os.path.exists(dst)) or os.path.getmtime(src) > os.path.getmtime(dst)


def build_c_oracle(force: bool = False) -> str:
    src = os.path.join(HERE, "balltree_oracle.c")
    out = os.path.join(HERE, "libballtree_oracle.so")
    if force or _newer(src, out):
        # -ffp-contract=off: the fp32 rotation must round after every multiply
        # and add exactly like the CUDA kernel's __fmul_rn/__fadd_rn sequence.
        cmd = [CC, "-O2", "-std=gnu11", "-fPIC", "-shared", "-ffp-contract=off",
               "-Wall", "-o", out, src, "-lm"]
        subprocess.check_call(cmd)
    return out


def ref_so_path() -> str | None:
    if not os.path.isdir(OUT_REF):
        return None
    for f in sorted(os.listdir(OUT_REF)):
        if f.startswith("balltree") and f.endswith(".so"):
            return os.path.join(OUT_REF, f)
    return None


def build_reference_balltree(force: bool = False) -> str | None:
    """Build the reference Cython extension into oracle/_ref/. Returns the path
    or None when /root/reference is absent (e.g. on the GPU box)."""
    existing = ref_so_path()
    pyx = os.path.join(REF_DIR, "balltree.pyx")
    if not os.path.exists(pyx):
        return existing
    if existing and not force and not _newer(pyx, existing):
        return existing
    import numpy

    os.makedirs(OUT_REF, exist_ok=True)
    gen = os.path.join(OUT_REF, "balltree_gen.cpp")
    subprocess.check_call([sys.executable, "-m", "cython", "--cplus", "-3",
                           "-o", gen, pyx])
    ext = sysconfig.get_config_var("EXT_SUFFIX")
    out = os.path.join(OUT_REF, "balltree" + ext)
    cmd = [CXX, "-O3", "-std=c++11", "-fopenmp", "-fPIC", "-shared", "-w",
           "-I", REF_DIR, "-I", numpy.get_include(),
           "-I", sysconfig.get_paths()["include"], "-o", out, gen]
    subprocess.check_call(cmd)
    os.remove(gen)
    return out


def build_all(force: bool = False) -> dict:
    return {"c_oracle": build_c_oracle(force),
            "reference_balltree": build_reference_balltree(force)}


if __name__ == "__main__":
    print(build_all(force="--force" in sys.argv))
