"""Kernel bindings: `_lib` (ctypes loader for libgmt_b200.so) and `local` (single-rank ops)."""
