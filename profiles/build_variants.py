"""Builds diagnostic variants of the library (extra -D flags) next to the product .so for in-box A/B runs
(profiles/ab_stage.py with RAC_LIB=<path>).  Usage: python profiles/build_variants.py name=-DFLAG1,-DFLAG2 ..."""
import importlib.util
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = os.path.join(ROOT, "end-to-end-rgb-pose-estimation-baseline_b200")
spec = importlib.util.spec_from_file_location("_rac_build", os.path.join(PKG, "build.py"))
b = importlib.util.module_from_spec(spec)
spec.loader.exec_module(b)
for arg in sys.argv[1:]:
    name, _, flags = arg.partition("=")
    out = os.path.join(PKG, "librac_b200_%s.so" % name)
    b.build_variant(out, [f for f in flags.split(",") if f])
    print(out)
