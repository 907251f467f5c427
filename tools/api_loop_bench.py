#!/usr/bin/env python
"""The reference's Python loop on device buffers — sample(selected masks); step(actions) — as two kernels
(k_sample + k_step_single<false>) and as one (cog_vec_step_sampled, k_step_single<true>); under ncu it exposes
the single-step kernels.   python tools/api_loop_bench.py [lib.so]      env: VB_ENVS, VB_ITERS, VB_DIFF, VB_MAXSTEPS, VB_MODE"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from gym_eldorado_b200 import capi, layout
L = capi.load(sys.argv[1]) if len(sys.argv) > 1 else capi.lib()
n = int(os.environ.get("VB_ENVS", "65536")); it = int(os.environ.get("VB_ITERS", "200"))
diff = int(os.environ.get("VB_DIFF", "1")); max_steps = int(os.environ.get("VB_MAXSTEPS", "100000"))
modes = os.environ.get("VB_MODE", "two,one").split(",")
name = os.path.basename(sys.argv[1]) if len(sys.argv) > 1 else "lib"
for mode in modes:
    v = capi.VecHandle(n, 0, 0, None, L); s = capi.SamplerHandle(n, 12345, 0, 0, None, L)
    v.reset(12345, 4, 3, diff, max_steps)
    mp, ap = v.buffer(layout.BUF_SELECTED_MASKS)[0], s.actions_ptr()
    if mode == "two":
        def go():
            s.sample_device(mp); v.step_device(ap)
    else:
        def go():
            v.step_sampled(s)
    for _ in range(10): go()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize(); e0.record()
    for _ in range(it): go()
    e1.record(); torch.cuda.synchronize()
    us = e0.elapsed_time(e1) * 1e3 / it
    try:
        last = "%.1f us" % (v.last_kernel_ms() * 1e3)
    except capi.CogError:
        last = "not timed"
    print("%-24s %-28s %8.1f us per sample+step  %.3e env-steps/s  (last kernel %s)" % (
        name, "k_sample + k_step_single" if mode == "two" else "cog_vec_step_sampled", us, n / us * 1e6, last))
    v.close(); s.close()
