"""Builds the host side in-tree with g++ (no CUDA sources here):
    graph-cut-ransac_b200/libgcr_host.so            the C++ surface (findRectifyingHomographySIFT_) over the C ABI
    graph-cut-ransac_b200/pygcransac.<abi>.so       the pybind11 module, same API as the reference's pygcransac
Both link libgcr_b200.so through an $ORIGIN rpath so that the package directory is self-contained."""
from __future__ import annotations

import os
import subprocess
import sys
import sysconfig

HERE = os.path.dirname(os.path.abspath(__file__))
PKG = os.path.dirname(os.path.dirname(HERE))


def _stale(target, sources):
    if not os.path.exists(target):
        return True
    t = os.path.getmtime(target)
    return any(os.path.getmtime(s) > t for s in sources)


def build(force: bool = False) -> str:
    import pybind11

    engine = os.path.join(PKG, "libgcr_b200.so")
    if not os.path.exists(engine):
        raise RuntimeError(f"{engine} must be built first (nvcc)")
    hdrs = [os.path.join(HERE, f) for f in os.listdir(HERE) if f.endswith(".hpp")]
    hdrs.append(os.path.join(os.path.dirname(PKG), "include", "gcr_b200.h"))
    # the system g++ by name: this image's $CXX wrapper (/opt/gcc) resolves -lstdc++ to a static archive, and a
    # module with its own private copy of libstdc++ crashes in iostream code once loaded next to the shared one
    cxx = os.environ.get("GCR_HOST_CXX", "g++")
    common = [cxx, "-std=c++17", "-O2", "-fPIC", "-shared", "-fvisibility=hidden", "-ffp-contract=off", "-Wall",
              "-Wno-unused-function"]
    link = ["-L" + PKG, "-Wl,-rpath,$ORIGIN", "-l:libgcr_b200.so"]

    host_lib = os.path.join(PKG, "libgcr_host.so")
    host_src = os.path.join(HERE, "gcr_host.cpp")
    if force or _stale(host_lib, [host_src, engine] + hdrs):
        subprocess.check_call(common + ["-fvisibility=default", host_src, "-o", host_lib] + link + ["-Wl,--no-undefined"])

    ext = sysconfig.get_config_var("EXT_SUFFIX")
    mod = os.path.join(PKG, "pygcransac" + ext)
    mod_src = os.path.join(HERE, "pygcransac_module.cpp")
    if force or _stale(mod, [mod_src, host_lib] + hdrs):
        subprocess.check_call(common + ["-I" + pybind11.get_include(), "-I" + sysconfig.get_paths()["include"], mod_src,
                                        "-o", mod, "-l:libgcr_host.so"] + link)
    return mod


if __name__ == "__main__":
    print(build(force="--force" in sys.argv))
