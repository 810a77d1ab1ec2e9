"""TEST INFRASTRUCTURE ONLY -- builds the reference's own Cython region filter.

Compiles /root/reference/cython_filter_new/filter_cython_new.pyx *where it lies*
(the source is never copied into this repo) into oracle/_ref/filter_cython_new*.so.
The reference's own setup.py (cython_filter_new/setup.py:1-10) uses distutils, which
is gone in Python 3.12, so this is the 5-line setuptools equivalent.

Only tests/, __graft_entry__ and bench.py's cpu_baseline / --impl reference legs may
import the result.  /root/reference does not exist on the GPU box: the .so built here
travels with the snapshot (oracle/_ref is git-ignored, not gpurun-ignored).
"""
import os
import shutil
import sys
import tempfile

REF_PYX = "/root/reference/cython_filter_new/filter_cython_new.pyx"
OUT_DIR = os.path.join(os.path.dirname(os.path.abspath(__file__)), "_ref")


def built_path():
    if not os.path.isdir(OUT_DIR):
        return None
    for f in os.listdir(OUT_DIR):
        if f.startswith("filter_cython_new") and f.endswith(".so"):
            return os.path.join(OUT_DIR, f)
    return None


def build(force=False):
    if built_path() and not force:
        return built_path()
    if not os.path.exists(REF_PYX):
        return None  # GPU box: only the prebuilt .so is available
    import numpy
    from Cython.Build import cythonize
    from setuptools import Extension
    from setuptools.dist import Distribution

    os.makedirs(OUT_DIR, exist_ok=True)
    work = tempfile.mkdtemp(prefix="e2sec_ref_")
    try:
        # cythonize writes its generated .c next to the build dir, not next to the .pyx
        ext = Extension("filter_cython_new", [REF_PYX], include_dirs=[numpy.get_include()],
                        define_macros=[("NPY_NO_DEPRECATED_API", "NPY_1_7_API_VERSION")])
        exts = cythonize([ext], language_level=3, build_dir=work, quiet=True)
        dist = Distribution({"ext_modules": exts})
        cmd = dist.get_command_obj("build_ext")
        cmd.build_lib = OUT_DIR
        cmd.build_temp = work
        cmd.ensure_finalized()
        cmd.run()
    finally:
        shutil.rmtree(work, ignore_errors=True)
    return built_path()


def load():
    """Import the compiled reference filter (module `filter_cython_new`)."""
    p = built_path() or build()
    if p is None:
        raise ImportError("oracle/_ref/filter_cython_new*.so not built and /root/reference absent")
    import importlib.util
    spec = importlib.util.spec_from_file_location("filter_cython_new", p)
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


if __name__ == "__main__":
    print(build(force="--force" in sys.argv))
