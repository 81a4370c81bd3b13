"""In-tree build of the native library (see build.py)."""
