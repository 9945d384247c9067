"""pll-modules_b200 — B200-native phylogenetic-likelihood hot path behind the libpll / pll-modules C API.

The product is the C-ABI shared library ``lib/libpll_b200.so`` (declared in ``include/pllb200.h``).
This Python module only locates/builds/loads it and exposes a few additive helpers; it contains no
arithmetic and no CPU fallback: if the library cannot be loaded, importing callers fail loudly.
"""
import ctypes
import os

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
LIB_PATH = os.path.join(HERE, "lib", "libpll_b200.so")
HARNESS_PATH = os.path.join(HERE, "lib", "libharness_b200.so")
INCLUDE_DIR = os.path.join(ROOT, "include")

_lib = None


class Stats(ctypes.Structure):
    _fields_ = [("kernel_launches", ctypes.c_ulonglong), ("h2d_bytes", ctypes.c_ulonglong),
                ("d2h_bytes", ctypes.c_ulonglong), ("newton_iterations", ctypes.c_ulonglong)]


def load():
    """dlopen libpll_b200.so (RTLD_LOCAL). Raises OSError if it has not been built."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise OSError("libpll_b200.so is not built: run python pll-modules_b200/build.py "
                          "(there is no CPU fallback)")
        lib = ctypes.CDLL(LIB_PATH, mode=ctypes.RTLD_LOCAL)
        lib.pllb200_version.restype = ctypes.c_char_p
        lib.pllb200_device_count.restype = ctypes.c_int
        lib.pllb200_set_option.argtypes = [ctypes.c_char_p, ctypes.c_int]
        lib.pllb200_set_option.restype = ctypes.c_int
        lib.pllb200_get_stats.argtypes = [ctypes.POINTER(Stats)]
        lib.pllb200_partition_set_stream.argtypes = [ctypes.c_void_p, ctypes.c_void_p]
        lib.pllb200_partition_synchronize.argtypes = [ctypes.c_void_p]
        lib.pllb200_set_device.argtypes = [ctypes.c_int]
        lib.pllb200_exchange_create.argtypes = [ctypes.c_void_p, ctypes.c_void_p]
        lib.pllb200_exchange_connect.argtypes = [ctypes.c_void_p, ctypes.c_int, ctypes.c_int, ctypes.c_void_p]
        lib.pllb200_exchange_ranks.argtypes = [ctypes.c_void_p]
        _lib = lib
    return _lib


def version():
    return load().pllb200_version().decode()


def device_count():
    return load().pllb200_device_count()


def stats():
    s = Stats()
    load().pllb200_get_stats(ctypes.byref(s))
    return {"kernel_launches": s.kernel_launches, "h2d_bytes": s.h2d_bytes, "d2h_bytes": s.d2h_bytes,
            "newton_iterations": s.newton_iterations}


def reset_stats():
    load().pllb200_reset_stats()


def set_option(name, value):
    return load().pllb200_set_option(name.encode(), int(value))


def last_eval_kernel_us():
    """Device time of the most recent logL / derivative kernel (option "eval_timing" must be 1), microseconds."""
    lib = load()
    lib.pllb200_last_eval_kernel_us.restype = ctypes.c_double
    return lib.pllb200_last_eval_kernel_us()


def exchange_create(partition_ptr):
    """CUDA IPC handle (64 bytes) of this rank's exchange buffer for the multi-GPU Newton kernel."""
    buf = ctypes.create_string_buffer(64)
    if not load().pllb200_exchange_create(partition_ptr, buf):
        raise RuntimeError("pllb200_exchange_create failed")
    return buf.raw


def exchange_connect(partition_ptr, rank, nranks, handles):
    """handles: the ranks' 64-byte handles concatenated in rank order."""
    assert len(handles) == 64 * nranks
    if not load().pllb200_exchange_connect(partition_ptr, rank, nranks, ctypes.c_char_p(handles)):
        raise RuntimeError("pllb200_exchange_connect failed")


def declared_symbols():
    """Function names declared in include/pllb200.h (used by the export test)."""
    import re
    text = open(os.path.join(INCLUDE_DIR, "pllb200.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    names = set(re.findall(r"\b(pll_[a-z0-9_]+|pllmod_[a-z0-9_]+|pllb200_[a-z0-9_]+)\s*\(", text))
    # typedef'd struct tags / types are not functions
    return sorted(n for n in names if not n.endswith("_t"))
