"""Development aid: build an A/B variant of the library with extra nvcc flags.

    python tools/build_variant.py NAME -DAKB_SOMETHING ...   ->  akarirender-1_b200/libakari_b200_NAME.so

Select it at run time with AKB_LIB=/root/repo/akarirender-1_b200/libakari_b200_NAME.so (abi.lib_path)."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import __graft_entry__ as g  # noqa: E402

name, extra = sys.argv[1], sys.argv[2:]
out = os.path.join(g.PKG_DIR, "libakari_b200_%s.so" % name)
print(g.build_library(force=True, out=out, extra_flags=extra))
