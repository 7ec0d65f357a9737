"""gfd-b200: B200-native backend for GPU Feature Discovery.

The product is native: `libgfd_b200.so` (C ABI, include/gfd_b200.h) and the `gpu-feature-discovery`
binary built from labeler/*.cpp.  This Python package is only the ctypes view of that C ABI that the
tests, bench.py and __graft_entry__.py use; it contains no label logic and no CPU fallback -- if the
shared library is missing, importing `binding` raises.

The directory name carries a hyphen (it mirrors the reference repo's name), so import it with
    importlib.import_module("gpu-feature-discovery_b200.binding")
"""
