#!/usr/bin/env python
"""Counts the memory and synchronisation instructions in the SASS of each all-reduce and
elementwise kernel of libb200validate.so (cuobjdump -sass; runs on the build box, no GPU).
This is the static evidence that the hand-written all-reduce kernels use multimem (NVSwitch)
instructions and system-scope peer accesses, and that the elementwise kernels issue 256-bit
global accesses:   python tools/sass_summary.py > profiles/r02_sass_summary.txt"""
import collections
import os
import re
import subprocess

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "nvidia-gpu-operator-k8s-cluster_b200", "lib", "libb200validate.so")
KEEP = re.compile(r"^(LDG|STG|LDGMC|STGMC|REDG|RED|ATOMG|ATOM|MEMBAR|CCTL|FENCE|ERRBAR|UBLKCP|SYNCS|BAR|LDS|STS|UCGABAR)")


def main():
    sass = subprocess.check_output(["cuobjdump", "-sass", LIB], text=True)
    fn, counts = None, collections.OrderedDict()
    for line in sass.splitlines():
        m = re.search(r"Function : (\S+)", line)
        if m:
            fn = subprocess.check_output(["c++filt", m.group(1)], text=True).strip()
            fn = re.sub(r"\(anonymous namespace\)::", "", fn)
            fn = re.sub(r"\(.*", "", fn).replace("void ", "")
            counts[fn] = collections.Counter()
            continue
        m = re.search(r"/\*[0-9a-f]{4}\*/\s+(?:@!?U?P\d+\s+)?([A-Z][A-Za-z0-9_.]+)", line)
        if fn and m and KEEP.match(m.group(1)):
            counts[fn][m.group(1)] += 1
    for fn, c in counts.items():
        if not re.search(r"allreduce|ew_tile_kernel<4, 8, 1, 1024|ew_tile_kernel<3, 8, 1, 1024", fn):
            continue
        print(fn)
        for op, k in sorted(c.items()):
            print("    %4d  %s" % (k, op))


if __name__ == "__main__":
    main()
