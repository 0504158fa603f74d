#!/usr/bin/env python
"""Tuning helper: build the engine with extra -D flags into build/variants/<name>.so (ships to the GPU box, not to git)
and time them on the box with   BRB_LIB=build/variants/<name>.so python bench.py --no-cpu-baseline.
usage: tools/variants.py name=-DVAR_X=1,-DVAR_Y=2 [name2=...]"""
import os, subprocess, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import __graft_entry__ as g
out = os.path.join(g.ROOT, "build", "variants"); os.makedirs(out, exist_ok=True)
procs = []
for arg in sys.argv[1:]:
    name, _, flags = arg.partition("=")
    cmd = [g.NVCC] + g.NVCC_FLAGS + [f for f in flags.split(",") if f] + ["-o", os.path.join(out, name + ".so"),
          os.path.join(g.CSRC, "engine.cu"), os.path.join(g.CSRC, "host_util.cpp")]
    procs.append((name, subprocess.Popen(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)))
for name, p in procs:
    o, _ = p.communicate()
    print(name, "ok" if p.returncode == 0 else "FAILED\n" + o[-2000:])
