#!/usr/bin/env python3
"""Development: role timeline of network_freq.cu (build with -DFN_TRACE: FN_TRACE=1 python rtx_nerf_b200/build.py)."""
import ctypes as C
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from rtx_nerf_b200 import _lib, api, synth  # noqa: E402

net = api.FrequencyNetwork(synth.mlp_weights(112, 128, 8, 16, seed=0))
rows = 148 * 256 * 12
s = torch.rand((rows, 5), device="cuda") * 2 - 1
rad = torch.empty((rows, 4), dtype=torch.float32, device="cuda")
for _ in range(3):
    net.forward(None, s, radiance=rad)
torch.cuda.synchronize()
buf = np.zeros(3 * 1024, np.int64)
lib = _lib.load()
lib.rtxnerf_debug_freqnet_trace.argtypes = [C.c_void_p]
assert lib.rtxnerf_debug_freqnet_trace(buf.ctypes.data) == 0
e0, e1, m = buf[:1024], buf[1024:2048], buf[2048:]
t0 = m[0]
P = 6   # pair index to print (steady state)
print("pair", P, "relative cycles")
eb = P * 16
mb = P * 36
for l in range(9):
    a0, i0, a1, i1 = m[mb + 4 * l: mb + 4 * l + 4] - t0
    line = "L%d  MMA: A0 seen %7d issued %7d | A1 seen %7d issued %7d" % (l, a0, i0, a1, i1)
    if l < 8:
        line += " || E0: D seen %7d pub %7d | E1: D seen %7d pub %7d" % (
            e0[eb + 2 * l] - t0, e0[eb + 1 + 2 * l] - t0, e1[eb + 2 * l] - t0, e1[eb + 1 + 2 * l] - t0)
    print(line)
