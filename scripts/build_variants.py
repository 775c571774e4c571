#!/usr/bin/env python
"""Build experimental variants of libpsla_frontend.so (psla_b200/csrc/variants/, git-ignored) in parallel.

    python scripts/build_variants.py name1:DEF1,DEF2 name2: ...
"""
import os
import sys
from concurrent.futures import ThreadPoolExecutor

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from psla_b200 import build  # noqa: E402


def one(spec):
    name, _, defs = spec.partition(":")  # name:DEF1,DEF2  -- an entry starting with '-' is passed to nvcc as a flag (use ';' for ',')
    items = [d for d in defs.split(",") if d]
    path = build.build_variant(name, [d for d in items if not d.startswith("-")],
                               extra_flags=[f.replace(";", ",") for d in items if d.startswith("-") for f in d.split(" ")])
    regs = [ln.strip() for ln in open(path + ".ptxas.txt") if "registers" in ln or "spill" in ln]
    return name, path, regs


if __name__ == "__main__":
    with ThreadPoolExecutor(max_workers=8) as ex:
        for name, path, regs in ex.map(one, sys.argv[1:]):
            spills = [r for r in regs if "spill" in r and "0 bytes spill stores" not in r]
            used = sorted({r.split("Used ")[1].split(" registers")[0] for r in regs if "Used" in r})
            print(f"{name}: {path} registers {used} spills {spills[:2]}")
