"""ctypes binding of the host-side harness in libmpas_b200.so (mesh, decomposition, JW init).

The arrays returned are numpy views of the library's own buffers, in the reference layout
(Fortran order: fastest index first, trailing garbage element), so `field("u@1")` has numpy
shape (nEdges+1, nVertLevels).
"""
import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libmpas_b200.so")
_lib = None


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise RuntimeError(
                f"{LIB_PATH} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
                "(there is no CPU fallback)")
        _lib = C.CDLL(LIB_PATH, mode=C.RTLD_GLOBAL)
        _lib.mpas_b200_last_error.restype = C.c_char_p
        _lib.mpas_host_case_create.restype = C.c_void_p
        _lib.mpas_host_case_create.argtypes = [C.c_int, C.c_int, C.c_int]
        _lib.mpas_host_case_destroy.argtypes = [C.c_void_p]
        _lib.mpas_host_case_dims.argtypes = [C.c_void_p] + [C.POINTER(C.c_int)] * 5
        _lib.mpas_host_case_permutation.argtypes = [C.c_void_p, C.c_int, C.c_void_p]
        _lib.mpas_host_case_partition.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_char_p]
        _lib.mpas_host_case_get_partition.argtypes = [C.c_void_p, C.c_void_p]
        _lib.mpas_host_case_build_block.argtypes = [C.c_void_p] + [C.c_int] * 6
        _lib.mpas_host_case_block.restype = C.c_void_p
        _lib.mpas_host_case_block.argtypes = [C.c_void_p, C.c_int]
        _lib.mpas_host_block_nfields.argtypes = [C.c_void_p]
        _lib.mpas_host_block_field_info.argtypes = [C.c_void_p, C.c_int, C.c_char_p, C.c_int, C.POINTER(C.c_int),
                                                    C.POINTER(C.c_int), C.POINTER(C.c_int), C.POINTER(C.c_longlong)]
        _lib.mpas_host_block_field_ptr.restype = C.c_void_p
        _lib.mpas_host_block_field_ptr.argtypes = [C.c_void_p, C.c_char_p]
        _lib.mpas_host_block_ncfg.argtypes = [C.c_void_p]
        _lib.mpas_host_block_cfg_info.argtypes = [C.c_void_p, C.c_int, C.c_char_p, C.c_int, C.POINTER(C.c_double)]
        _lib.mpas_host_block_nexch.argtypes = [C.c_void_p, C.c_int, C.c_int]
        _lib.mpas_host_block_exch_info.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int] + [C.POINTER(C.c_int)] * 3
        _lib.mpas_host_block_exch_lists.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p]
    return _lib


def last_error():
    return lib().mpas_b200_last_error().decode()


class Block:
    """One decomposed block: local mesh + pool of Registry-named fields + exchange lists."""

    def __init__(self, case, handle, iblock):
        self.case = case  # keeps the owning HostCase alive
        self.h = handle
        self.id = iblock
        self._info = None

    def _load_info(self):
        L = lib()
        self._info = {}
        name = C.create_string_buffer(128)
        kind, loc, nd = C.c_int(), C.c_int(), C.c_int()
        dims = (C.c_longlong * 8)()
        for i in range(L.mpas_host_block_nfields(self.h)):
            L.mpas_host_block_field_info(self.h, i, name, 128, C.byref(kind), C.byref(loc), C.byref(nd), dims)
            self._info[name.value.decode()] = (kind.value, loc.value, tuple(dims[k] for k in range(nd.value)))

    def names(self):
        if self._info is None:
            self._load_info()
        return list(self._info.keys())

    def info(self, name):
        if self._info is None or name not in self._info:
            self._load_info()
        return self._info[name]

    def field(self, name):
        kind, loc, dims = self.info(name)
        ptr = lib().mpas_host_block_field_ptr(self.h, name.encode())
        if not ptr:
            raise KeyError(name)
        n = int(np.prod(dims))
        ct = C.c_double if kind == 0 else C.c_int32
        arr = np.ctypeslib.as_array(C.cast(ptr, C.POINTER(ct)), shape=(n,))
        return arr.reshape(tuple(reversed(dims)))

    def cfg(self):
        L = lib()
        out = {}
        name = C.create_string_buffer(128)
        val = C.c_double()
        for i in range(L.mpas_host_block_ncfg(self.h)):
            L.mpas_host_block_cfg_info(self.h, i, name, 128, C.byref(val))
            out[name.value.decode()] = val.value
        return out

    def exchange_lists(self, kind, which):
        """kind: 'cell'|'edge'|'vertex'; which: 'send'|'recv' -> list of dicts in (endPoint, layer) order."""
        L = lib()
        k = {"cell": 0, "edge": 1, "vertex": 2}[kind]
        w = {"send": 0, "recv": 1}[which]
        out = []
        ep, hl, nl = C.c_int(), C.c_int(), C.c_int()
        for i in range(L.mpas_host_block_nexch(self.h, k, w)):
            L.mpas_host_block_exch_info(self.h, k, w, i, C.byref(ep), C.byref(hl), C.byref(nl))
            src = np.zeros(nl.value, dtype=np.int32)
            dst = np.zeros(nl.value, dtype=np.int32)
            L.mpas_host_block_exch_lists(self.h, k, w, i, src.ctypes.data, dst.ctypes.data)
            out.append(dict(endPoint=ep.value, layer=hl.value, src=src, dest=dst))
        return out


class HostCase:
    """Global icosahedral mesh + partition + per-block JW initial state."""

    def __init__(self, level, lloyd_iters=0, sfc=True):
        self.h = lib().mpas_host_case_create(level, lloyd_iters, 1 if sfc else 0)
        if not self.h:
            raise RuntimeError(last_error())
        d = [C.c_int() for _ in range(5)]
        lib().mpas_host_case_dims(self.h, *[C.byref(x) for x in d])
        self.nCells, self.nEdges, self.nVertices, self.maxEdges, self.maxEdges2 = [x.value for x in d]
        self.sfc = sfc
        self.nBlocks = 1

    def __del__(self):
        try:
            if self.h:
                lib().mpas_host_case_destroy(self.h)
                self.h = None
        except Exception:
            pass

    def permutation(self, kind):
        n = {"cell": self.nCells, "edge": self.nEdges, "vertex": self.nVertices}[kind]
        out = np.zeros(n, dtype=np.int32)
        if lib().mpas_host_case_permutation(self.h, {"cell": 0, "edge": 1, "vertex": 2}[kind], out.ctypes.data):
            raise RuntimeError(last_error())
        return out

    def partition(self, nblocks, method="metis", prefix=""):
        m = {"metis": 0, "sfc": 1, "file": 2}[method]
        if lib().mpas_host_case_partition(self.h, nblocks, m, prefix.encode()):
            raise RuntimeError(last_error())
        self.nBlocks = nblocks
        out = np.zeros(self.nCells, dtype=np.int32)
        lib().mpas_host_case_get_partition(self.h, out.ctypes.data)
        return out

    def build_block(self, iblock=0, nVertLevels=26, init_case=2, num_scalars=1, full_fields=True, keep_zb=False):
        rc = lib().mpas_host_case_build_block(self.h, iblock, nVertLevels, init_case, num_scalars, 1 if full_fields else 0,
                                              1 if keep_zb else 0)
        if rc:
            raise RuntimeError(last_error())
        return Block(self, lib().mpas_host_case_block(self.h, iblock), iblock)
