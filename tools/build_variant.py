#!/usr/bin/env python
"""Build a variant of the CUDA library with extra -D macros next to the product library (kernel A/B measurements):
    python tools/build_variant.py NAME MACRO[=VALUE] ...   ->  distributed-llm-scheduler_b200/variants/libdls_NAME.so
Run anything against it with DLS_B200_LIB=<that path>. Variant libraries are build products (git-ignored)."""
import os
import sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from dls_b200 import _native as N

name, defines = sys.argv[1], sys.argv[2:]
d = os.path.join(N.PKG_DIR, "variants")
os.makedirs(d, exist_ok=True)
print(N.build(force=True, defines=defines, out=os.path.join(d, f"libdls_{name}.so")))
