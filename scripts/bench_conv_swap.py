"""Timing experiment (profile-hooks library): conv_tc3 mainloop with the operands swapped (weights = A, 256 pixels = N),
epilogue off, against the shipped operand order.  OF_PROFILE_LIB=1 python scripts/bench_conv_swap.py"""
import os, sys, subprocess
here = os.path.dirname(os.path.abspath(__file__))
for dbg in ("2", "130"):
    env = dict(os.environ, OF_PROFILE_LIB="1", OF_T3_DBG=dbg, OF_CONV_SHAPES="pow2")
    print("== OF_T3_DBG=%s (%s)" % (dbg, {"2": "mainloop only, shipped order", "130": "mainloop only, swapped operands N=256", "0": "full kernel"}[dbg]), flush=True)
    subprocess.run([sys.executable, os.path.join(here, "bench_conv_layer.py")], env=env)
