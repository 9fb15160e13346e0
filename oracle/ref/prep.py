#!/usr/bin/env python
"""TEST INFRASTRUCTURE.  Rewrites the OpenCL vector-literal syntax of the reference's clcode.cl
into C++ constructor calls so that it compiles with g++ over clshim.h:

    (float2)(a, b)  ->  float2(a, b)        (same for float3/4/8, int2/3, uchar3/8)

Nothing else is touched.  Input is read where it lies (/root/reference), output goes to
oracle/_ref/ (git-ignored).  usage: prep.py <clcode.cl> <out.inc>"""
import re
import sys

src = open(sys.argv[1], encoding="utf-8", errors="replace").read()
src = src.replace("\r\n", "\n")
out = re.sub(r"\(\s*(float[2348]|int[23]|uchar[38])\s*\)\s*\(", r"\1(", src)
open(sys.argv[2], "w").write(out)
print(f"prep: {len(src.splitlines())} lines, {len(re.findall(r'(float[2348]|int[23]|uchar[38])\(', out))} vector constructors")
