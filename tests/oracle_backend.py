"""Binds the CPU oracle (oracle/libb2oracle.so) to the api.World wrapper. TEST INFRASTRUCTURE ONLY."""
import ctypes as C
import os
import subprocess
from box2d_b200 import capi, api

_ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
_LIB = None


def oracle_lib():
    global _LIB
    if _LIB is None:
        d = os.path.join(_ROOT, "oracle")
        subprocess.check_call(["make", "-s", "-C", d])
        _LIB = C.CDLL(os.path.join(d, "libb2oracle.so"))
    return _LIB


def OracleBackend():
    lib = oracle_lib()
    fns = capi.bind(lib, "b2o_")
    lib.b2o_world_create.argtypes = [C.POINTER(capi.WorldDef), C.POINTER(C.c_void_p)]
    lib.b2o_last_error.argtypes = [C.c_void_p]
    lib.b2o_last_error.restype = C.c_char_p
    lib.b2o_world_get_new_pairs.argtypes = [C.c_void_p, C.POINTER(C.c_uint64), C.c_int32, C.POINTER(C.c_int32)]
    lib.b2o_world_get_tree_ranks.argtypes = [C.c_void_p, C.c_int32, C.POINTER(C.c_int32), C.c_int32, C.POINTER(C.c_int32)]

    def create(wdef):
        h = C.c_void_p()
        lib.b2o_world_create(C.byref(wdef), C.byref(h))
        return h

    return api.Backend(fns, create, lambda h: lib.b2o_last_error(h).decode(), "oracle")
