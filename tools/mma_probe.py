#!/usr/bin/env python
"""Clocks per MMA of the instruction mixes the streaming CRP kernels issue (acoss_debug_mma_probe, one SM, zero operands)."""
import ctypes, os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import acoss_b200
ctx = acoss_b200.Context(0)
names = ["14 TF32 (hi.hi)", "41 TF32 (three products)", "28 mixed (14 TF32 + 14 bf16)", "14 bf16, K chunks adjacent", "14 bf16, K chunks in two planes"]
n_mma = [14, 41, 28, 14, 14]
for reps in (8, 64):
    out = np.zeros(5, np.int64)
    ctx._ck(ctx.lib.acoss_debug_mma_probe(ctx.h, reps, out.ctypes.data_as(ctypes.POINTER(ctypes.c_int64))))
    for nm, k, t in zip(names, n_mma, out):
        print("reps %3d  %-34s %8d clk  = %6.1f clk per MMA" % (reps, nm, t, t / float(reps * k)))
