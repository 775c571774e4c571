#!/usr/bin/env python
"""Static SASS instruction mix of one kernel of a built library: python scripts/sass_mix.py lib.so [substring-of-mangled-name]"""
import re
import subprocess
import sys
from collections import Counter

lib = sys.argv[1]
key = sys.argv[2] if len(sys.argv) > 2 else "fbank_stream_kernelILb0ELb1ELb1EfLb0E"
txt = subprocess.run(["cuobjdump", "-sass", lib], capture_output=True, text=True).stdout
for f in re.split(r"\n\s*Function : ", txt)[1:]:
    name = f.split("\n")[0]
    if key in name:
        ins = re.findall(r"/\*[0-9a-f]{4,6}\*/\s+(?:@!?U?P\d\s+)?([A-Z0-9_]+)", f)
        c = Counter(ins)
        print(lib.split("/")[-1], "total", len(ins), " ".join(f"{k}:{v}" for k, v in c.most_common(45)))
