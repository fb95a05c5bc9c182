"""Native-library plumbing: in-tree build (build.py) and ctypes bindings (cuda_lib / cpu_lib)."""
