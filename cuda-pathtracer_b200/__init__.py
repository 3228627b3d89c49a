"""B200-native path-tracing core: ctypes host binding over libptx.so (see binding.py, include/ptx.h)."""
from .binding import *  # noqa: F401,F403
from .binding import lib, EXPORTS, LIB_PATH  # noqa: F401
