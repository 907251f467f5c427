"""CPU: every kernel and host entry point of the library once through AddressSanitizer + UBSan.

compute-sanitizer is not available on the GPU pool; the warp emulator (tests/emu — test infrastructure) runs the
SAME csrc/*.cu[h] sources on the host with "device" memory on the host heap, so an out-of-bounds access, a
misaligned vector access or signed overflow in the kernels' logic is an ASan / UBSan report here.  The workload
(tools/sanitize_workload.py) has short episodes — board generation in every kernel that inlines it —, merged
masks, the rollout ring, compact records and all host-I/O forms (incl. the no-allocation entry point)."""
import os
import subprocess
import sys

import pytest

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
LIB = os.path.join(HERE, "emu", "libcog_emu_asan.so")

CHILD = r'''
import os, sys
sys.path.insert(0, %r)
from gym_eldorado_b200 import capi
emu = capi.load(%r)
capi.lib = lambda: emu
os.environ["VB_ENVS"] = "16"
path = os.path.join(%r, "tools", "sanitize_workload.py")
exec(compile(open(path).read(), path, "exec"), {"__file__": path, "__name__": "__main__"})
''' % (ROOT, LIB, ROOT)


def _runtime(name):
    path = subprocess.run(["g++", "-print-file-name=" + name], capture_output=True, text=True).stdout.strip()
    return path if os.path.isabs(path) and os.path.exists(path) else None


def test_emulated_kernels_under_asan_ubsan():
    asan, ubsan = _runtime("libasan.so"), _runtime("libubsan.so")
    if not asan or not ubsan:
        pytest.skip("g++ has no sanitizer runtimes here")
    emu = os.path.join(HERE, "emu")
    csrc = os.path.join(ROOT, "gym_eldorado_b200", "csrc")
    sources = [os.path.join(emu, f) for f in ("emu_lib.cpp", "cuda_emu.h")] + \
              [os.path.join(csrc, f) for f in os.listdir(csrc) if f.endswith((".cu", ".cuh", ".h"))]
    if not os.path.exists(LIB) or any(os.path.getmtime(f) > os.path.getmtime(LIB) for f in sources):
        subprocess.check_call(["g++", "-std=c++17", "-O1", "-g1", "-fPIC", "-shared", "-fvisibility=hidden",
                               "-fno-omit-frame-pointer", "-fsanitize=address,undefined",
                               "-fno-sanitize-recover=undefined", "-I" + emu, "-I" + os.path.join(ROOT, "include"),
                               "-I" + csrc, os.path.join(emu, "emu_lib.cpp"), "-o", LIB])
    env = dict(os.environ, LD_PRELOAD=asan + " " + ubsan,
               # the emulator switches stacks by hand (one fiber per lane): no fake stacks; CPython "leaks" by design
               ASAN_OPTIONS="detect_leaks=0:detect_stack_use_after_return=0:abort_on_error=0",
               UBSAN_OPTIONS="print_stacktrace=1")
    r = subprocess.run([sys.executable, "-c", CHILD], env=env, capture_output=True, text=True, timeout=1500)
    assert r.returncode == 0, (r.stdout + r.stderr)[-4000:]
    assert "ERROR: AddressSanitizer" not in r.stderr and "runtime error:" not in r.stderr, r.stderr[-4000:]
    assert r.stdout.count(" ok ") == 3, r.stdout
