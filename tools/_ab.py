"""A/B aid for the timing tools: BQA_LIB_PATH=<other libbqa.so> makes a TOOL time another build of the
library on the same box.  The product loader (_lib.py) always loads the in-tree libbqa.so; only the tools
that import this module honour the variable."""
import os


def maybe_swap_library(pkg):
    path = os.environ.get("BQA_LIB_PATH")
    if path:
        L = pkg._lib
        L.lib = L._load(path)  # same ABI check, same signatures; every caller looks `L.lib` up at call time
        L.LIB_PATH = path
    return pkg
