"""TEST INFRASTRUCTURE ONLY - compiles the reference's own native code for the checker, from the sources where they lie.

    python -m oracle.build_ref            # needs /root/reference; writes ONLY into oracle/_ref/ (git-ignored)

The one native component of the reference is the Cython monotonic alignment search
(/root/reference/matcha/utils/monotonic_align/core.pyx).  Recipe: cython translates the .pyx to C (output file under
oracle/_ref/, nothing is written next to the source), gcc builds the extension module.  The reference's own build
system (its setup.py) is not run.  tests/test_mas.py loads the module with `load_ref_mas()` to pin oracle/mas.py and to
generate tests/golden/mas_*.npz (oracle/make_golden.py mas); on the GPU box the prebuilt .so travels with the snapshot.
"""
import importlib.util
import os
import subprocess
import sys
import sysconfig

HERE = os.path.dirname(os.path.abspath(__file__))
OUT = os.path.join(HERE, "_ref")
PYX = "/root/reference/matcha/utils/monotonic_align/core.pyx"


def so_path():
    return os.path.join(OUT, "core" + sysconfig.get_config_var("EXT_SUFFIX"))


def build():
    """Returns the path of the built module, or None when the reference sources are not present."""
    if not os.path.exists(PYX):
        return so_path() if os.path.exists(so_path()) else None
    import numpy
    os.makedirs(OUT, exist_ok=True)
    c_file = os.path.join(OUT, "core.c")
    subprocess.check_call([sys.executable, "-m", "cython", "-3", PYX, "-o", c_file])
    subprocess.check_call(["gcc", "-O2", "-shared", "-fPIC", "-fopenmp", "-DNPY_NO_DEPRECATED_API=NPY_1_7_API_VERSION",
                           "-I" + sysconfig.get_paths()["include"], "-I" + numpy.get_include(), c_file, "-o", so_path()])
    return so_path()


def load_ref_mas():
    """The compiled reference module (`maximum_path_c(paths, values, t_xs, t_ys)`), or None if it has not been built."""
    if not os.path.exists(so_path()):
        return None
    spec = importlib.util.spec_from_file_location("core", so_path())
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


if __name__ == "__main__":
    print(build())
