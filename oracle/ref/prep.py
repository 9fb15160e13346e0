#!/usr/bin/env python
"""TEST INFRASTRUCTURE.  Rewrites the OpenCL vector-literal syntax of the reference's clcode.cl
into C++ constructor calls so that it compiles with g++ over clshim.h:

    (float2)(a, b)  ->  float2(a, b)        (same for float3/4/8, int2/3, uchar3/8)

and gives the ninth store of edge_compute_alternative a home: its `float3 color[8]` is indexed 0..8 (clcode.cl:169-184;
the ninth value is never read), which as C++ would write past the array:

    float3 color[8];  ->  float3 color[9];

Nothing else is touched.  Input is read where it lies (/root/reference), output goes to
oracle/_ref/ (git-ignored).  usage: prep.py <clcode.cl> <out.inc>"""
import re
import sys

src = open(sys.argv[1], encoding="utf-8", errors="replace").read()
src = src.replace("\r\n", "\n")
out = re.sub(r"\(\s*(float[2348]|int[23]|uchar[38])\s*\)\s*\(", r"\1(", src)
out, n_color = re.subn(r"float3\s+color\[8\];", "float3 color[9];", out)
assert n_color == 1, "edge_compute_alternative: expected exactly one `float3 color[8];`"
open(sys.argv[2], "w").write(out)
print(f"prep: {len(src.splitlines())} lines, {len(re.findall(r'(float[2348]|int[23]|uchar[38])\(', out))} vector constructors")
