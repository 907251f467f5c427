"""Builds the pybind11 host module `city_of_gold` (csrc/pybind_module.cpp) in-tree and links
it against libcog_b200.so (rpath $ORIGIN).  g++ only: the module holds no CUDA code."""
import os
import subprocess
import sys
import sysconfig

PKG = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(PKG)
SRC = os.path.join(PKG, "csrc", "pybind_module.cpp")
EXT = sysconfig.get_config_var("EXT_SUFFIX")
TARGET = os.path.join(PKG, "city_of_gold" + EXT)


def compile_module(target, lib_dir, lib_name, module_name, extra=()):
    import pybind11
    cxx = "/usr/bin/g++" if os.path.exists("/usr/bin/g++") else "g++"
    cmd = [cxx, "-std=c++17", "-O2", "-fPIC", "-shared", "-fvisibility=hidden",
           "-DCOG_MODULE_NAME=" + module_name, *extra,
           "-I" + os.path.join(ROOT, "include"), "-I" + pybind11.get_include(),
           "-I" + sysconfig.get_paths()["include"], SRC, "-o", target,
           "-L" + lib_dir, "-l:" + lib_name, "-Wl,-rpath,$ORIGIN"]
    subprocess.check_call(cmd)


def build(force=False):
    deps = [SRC, os.path.join(ROOT, "include", "cog_b200.h")]
    if force or not os.path.exists(TARGET) or any(os.path.getmtime(d) > os.path.getmtime(TARGET) for d in deps):
        compile_module(TARGET, PKG, "libcog_b200.so", "city_of_gold")
    return TARGET


if __name__ == "__main__":
    build("--force" in sys.argv)
