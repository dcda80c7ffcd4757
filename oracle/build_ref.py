"""Build the UNMODIFIED reference calculator into oracle/_ref/ -- TEST INFRASTRUCTURE ONLY.

Recipe: /root/reference is read-only and the build writes
beside the sources, so the tree is copied to a scratch directory, BitArray's
``bit_array.c`` is compiled with gcc, the six Cython modules on (or beside) the hot
path are cythonized with the reference's own flags (setup.py:113: ``-O3 -ffast-math``,
no ``-mpopcnt``), and the resulting package -- the reference's .py files plus the
built extension modules -- is installed into oracle/_ref/ (git-ignored; it travels to
the GPU box with the snapshot).  No reference source is copied into tracked paths.

``core/readlen.pyx`` is not built (cimports pysam, absent from this image).
``PyMaSC.reader.bigwig`` imports a module named ``pyBigWig``; oracle/ref_stubs/pyBigWig.py
(our own code) stands in and serves intervals from the fixtures' .bedGraph twin.

Usage:  python oracle/build_ref.py [--popcnt]     (--popcnt -> oracle/_ref_popcnt, "-O3 -mpopcnt")
"""
import glob
import os
import shutil
import subprocess
import sys
import tempfile

HERE = os.path.dirname(os.path.abspath(__file__))
REFERENCE = os.environ.get("PYMASC_REFERENCE", "/root/reference")

SETUP = r'''
import numpy
from setuptools import setup, Extension
from Cython.Build import cythonize
FLAGS = %(flags)r
BA = ["external/BitArray", "external"]
NP = [numpy.get_include()]
def ext(name, src, inc, link=()):
    return Extension(name, [src], include_dirs=inc, extra_link_args=list(link),
                     extra_compile_args=FLAGS)
exts = [
    ext("PyMaSC.reader.bigwig", "PyMaSC/reader/bigwig.pyx", NP),
    ext("PyMaSC.core.mappability", "PyMaSC/core/mappability.pyx", NP),
    ext("PyMaSC.core.successive.ncc", "PyMaSC/core/successive/ncc.pyx", NP),
    ext("PyMaSC.core.successive.mscc", "PyMaSC/core/successive/mscc.pyx", NP),
    ext("PyMaSC.core.bitarray.bitarray", "PyMaSC/core/bitarray/bitarray.pyx", BA + NP,
        ["external/BitArray/bit_array.o"]),
    ext("PyMaSC.core.bitarray.mscc", "PyMaSC/core/bitarray/mscc.pyx", BA + NP,
        ["external/BitArray/bit_array.o"]),
]
setup(name="pymasc_ref_oracle",
      ext_modules=cythonize(exts, compiler_directives={"language_level": 3}, quiet=True),
      script_args=["-q", "build_ext", "--inplace"])
'''


def available():
    return os.path.isdir(os.path.join(REFERENCE, "PyMaSC"))


def build(popcnt=False, force=False):
    out = os.path.join(HERE, "_ref_popcnt" if popcnt else "_ref")
    marker = os.path.join(out, "PyMaSC", "core", "bitarray")
    if not force and glob.glob(os.path.join(marker, "mscc*.so")):
        return out
    if not available():
        raise RuntimeError("reference checkout not found at %s" % REFERENCE)
    flags = ["-O3", "-ffast-math"] + (["-mpopcnt"] if popcnt else [])
    work = tempfile.mkdtemp(prefix="pymasc_refbuild_")
    try:
        src = os.path.join(work, "src")
        shutil.copytree(REFERENCE, src, symlinks=True,
                        ignore=shutil.ignore_patterns(".git", "*.so", "*.o", "*.a", "build"))
        subprocess.check_call("chmod -R u+w .", shell=True, cwd=src)
        ba = os.path.join(src, "external", "BitArray")
        # external/BitArray/Makefile:9-19,27-28 (gcc: -O4 == -O3; -fPIC object)
        subprocess.check_call(["gcc", "-Wall", "-Wextra", "-I.", "-O3", "-fPIC"] +
                              (["-mpopcnt"] if popcnt else []) +
                              ["-c", "bit_array.c", "-o", "bit_array.o"], cwd=ba)
        with open(os.path.join(src, "_oracle_setup.py"), "w") as fh:
            fh.write(SETUP % {"flags": flags})
        subprocess.check_call([sys.executable, "_oracle_setup.py"], cwd=src)
        if os.path.isdir(out):
            shutil.rmtree(out)
        shutil.copytree(os.path.join(src, "PyMaSC"), os.path.join(out, "PyMaSC"),
                        ignore=shutil.ignore_patterns("*.pyx", "*.pxd", "*.c", "*.cpp",
                                                      "__pycache__"))
        with open(os.path.join(out, "BUILD_INFO.txt"), "w") as fh:
            fh.write("built from %s with flags %s\n" % (REFERENCE, " ".join(flags)))
    finally:
        shutil.rmtree(work, ignore_errors=True)
    return out


if __name__ == "__main__":
    p = build(popcnt="--popcnt" in sys.argv, force="--force" in sys.argv)
    print(p)
