"""ctypes binding of the host-side harness (libmpas_host.so: mesh, decomposition, JW init, files -- no CUDA) and of the
device-resident dycore (libmpas_b200.so).  hostlib() never loads the CUDA library: the CPU oracle and the reference arm
of bench.py use it alone.

The arrays returned are numpy views of the library's own buffers, in the reference layout
(Fortran order: fastest index first, trailing garbage element), so `field("u@1")` has numpy
shape (nEdges+1, nVertLevels).
"""
import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
# MPAS_B200_PRECISION=single selects the PRECISION=single build (RKIND = 4 bytes, mpas_kind_types.F:25-29): the same
# sources compiled with -DSINGLE_PRECISION.  The host harness inside it stays double; real fields are converted on
# upload (dycore.cu: add_field) and by Dycore.field / set_field below.  One precision per process.
SINGLE = os.environ.get("MPAS_B200_PRECISION", "double").lower() in ("single", "sp", "4")
LIB_PATH = os.path.join(_HERE, "libmpas_b200_sp.so" if SINGLE else "libmpas_b200.so")
HOSTLIB_PATH = os.path.join(_HERE, "libmpas_host.so")
_lib = None
_hostlib = None


def hostlib():
    """the host harness library (no CUDA): mesh generator, partition, block decomposition, JW state, netCDF files"""
    global _hostlib
    if _hostlib is None:
        if not os.path.exists(HOSTLIB_PATH):
            raise RuntimeError(f"{HOSTLIB_PATH} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'`")
        _hostlib = C.CDLL(HOSTLIB_PATH, mode=C.RTLD_GLOBAL)
        _hostlib.mpas_host_last_error.restype = C.c_char_p
        _hostlib.mpas_host_case_create.restype = C.c_void_p
        _hostlib.mpas_host_case_create.argtypes = [C.c_int, C.c_int, C.c_int]
        _hostlib.mpas_host_case_create_planar.restype = C.c_void_p
        _hostlib.mpas_host_case_create_planar.argtypes = [C.c_int, C.c_int, C.c_double, C.c_int]
        _hostlib.mpas_host_case_destroy.argtypes = [C.c_void_p]
        _hostlib.mpas_host_case_dims.argtypes = [C.c_void_p] + [C.POINTER(C.c_int)] * 5
        _hostlib.mpas_host_case_permutation.argtypes = [C.c_void_p, C.c_int, C.c_void_p]
        _hostlib.mpas_host_case_partition.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_char_p]
        _hostlib.mpas_host_case_get_partition.argtypes = [C.c_void_p, C.c_void_p]
        _hostlib.mpas_host_case_build_block.argtypes = [C.c_void_p] + [C.c_int] * 6
        _hostlib.mpas_host_case_block.restype = C.c_void_p
        _hostlib.mpas_host_case_block.argtypes = [C.c_void_p, C.c_int]
        _hostlib.mpas_host_block_nfields.argtypes = [C.c_void_p]
        _hostlib.mpas_host_block_field_info.argtypes = [C.c_void_p, C.c_int, C.c_char_p, C.c_int, C.POINTER(C.c_int),
                                                    C.POINTER(C.c_int), C.POINTER(C.c_int), C.POINTER(C.c_longlong)]
        _hostlib.mpas_host_block_field_ptr.restype = C.c_void_p
        _hostlib.mpas_host_block_field_ptr.argtypes = [C.c_void_p, C.c_char_p]
        _hostlib.mpas_host_block_field_is_lean.argtypes = [C.c_void_p, C.c_char_p]
        _hostlib.mpas_host_block_ncfg.argtypes = [C.c_void_p]
        _hostlib.mpas_host_block_cfg_info.argtypes = [C.c_void_p, C.c_int, C.c_char_p, C.c_int, C.POINTER(C.c_double)]
        _hostlib.mpas_host_block_nexch.argtypes = [C.c_void_p, C.c_int, C.c_int]
        _hostlib.mpas_host_block_exch_info.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int] + [C.POINTER(C.c_int)] * 3
        _hostlib.mpas_host_block_exch_lists.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p]
        _hostlib.mpas_host_case_write_grid.argtypes = [C.c_void_p, C.c_char_p, C.c_int]
        _hostlib.mpas_host_case_create_from_grid.restype = C.c_void_p
        _hostlib.mpas_host_case_create_from_grid.argtypes = [C.c_char_p, C.c_int]
        _hostlib.mpas_host_case_create_restart.argtypes = [C.c_void_p, C.c_int, C.c_char_p, C.c_int, C.c_char_p, C.c_char_p]
        _hostlib.mpas_host_block_write_restart.argtypes = [C.c_void_p, C.c_char_p, C.c_char_p]
        _hostlib.mpas_host_block_read_restart.argtypes = [C.c_void_p, C.c_char_p]
        _hostlib.mpas_nc_open.restype = C.c_void_p
        _hostlib.mpas_nc_open.argtypes = [C.c_char_p]
        _hostlib.mpas_nc_create.restype = C.c_void_p
        _hostlib.mpas_nc_create.argtypes = [C.c_char_p, C.c_int]
        _hostlib.mpas_nc_close.argtypes = [C.c_void_p]
        _hostlib.mpas_nc_inq.argtypes = [C.c_void_p] + [C.POINTER(C.c_int)] * 4 + [C.POINTER(C.c_longlong)]
        _hostlib.mpas_nc_dim_info.argtypes = [C.c_void_p, C.c_int, C.c_char_p, C.c_int, C.POINTER(C.c_longlong)]
        _hostlib.mpas_nc_var_info.argtypes = [C.c_void_p, C.c_int, C.c_char_p, C.c_int, C.POINTER(C.c_int), C.POINTER(C.c_int), C.c_void_p,
                                          C.POINTER(C.c_longlong)]
        _hostlib.mpas_nc_gatt_text.argtypes = [C.c_void_p, C.c_char_p, C.c_char_p, C.c_int]
        _hostlib.mpas_nc_get_double.argtypes = [C.c_void_p, C.c_char_p, C.c_void_p]
        _hostlib.mpas_nc_get_int.argtypes = [C.c_void_p, C.c_char_p, C.c_void_p]
        _hostlib.mpas_nc_def_dim.argtypes = [C.c_void_p, C.c_char_p, C.c_longlong]
        _hostlib.mpas_nc_def_var.argtypes = [C.c_void_p, C.c_char_p, C.c_int, C.c_int, C.c_void_p]
        _hostlib.mpas_nc_put_gatt_text.argtypes = [C.c_void_p, C.c_char_p, C.c_char_p]
        _hostlib.mpas_nc_enddef.argtypes = [C.c_void_p, C.c_longlong]
        _hostlib.mpas_nc_put_double.argtypes = [C.c_void_p, C.c_int, C.c_longlong, C.c_longlong, C.c_longlong, C.c_void_p]
        _hostlib.mpas_nc_put_int.argtypes = [C.c_void_p, C.c_int, C.c_longlong, C.c_longlong, C.c_longlong, C.c_void_p]
    return _hostlib


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise RuntimeError(
                f"{LIB_PATH} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
                "(there is no CPU fallback)")
        _lib = C.CDLL(LIB_PATH, mode=C.RTLD_GLOBAL)
        if _lib.mpas_b200_real_size() != (4 if SINGLE else 8):
            raise RuntimeError(f"{LIB_PATH}: sizeof(mpas_real) = {_lib.mpas_b200_real_size()} does not match MPAS_B200_PRECISION")
        _lib.mpas_b200_last_error.restype = C.c_char_p
        _lib.mpas_b200_dycore_field_is_scratch_view.argtypes = [C.c_void_p, C.c_int, C.c_char_p]
        _lib.mpas_b200_dycore_device_bytes.restype = C.c_longlong
        _lib.mpas_b200_dycore_device_bytes.argtypes = [C.c_void_p]
        _lib.mpas_b200_dycore_checksum.argtypes = [C.c_void_p, C.c_char_p, C.POINTER(C.c_ulonglong)]
        # device-resident dycore (include/mpas_b200.h, layer 2) and registry (layer 1 lifecycle)
        _lib.mpas_b200_set_device.argtypes = [C.c_int]
        _lib.mpas_b200_register_field.argtypes = [C.c_void_p, C.c_size_t]
        _lib.mpas_b200_unregister_field.argtypes = [C.c_void_p]
        _lib.mpas_b200_upload.argtypes = [C.c_void_p]
        _lib.mpas_b200_download.argtypes = [C.c_void_p]
        _lib.mpas_b200_device_ptr.restype = C.c_void_p
        _lib.mpas_b200_device_ptr.argtypes = [C.c_void_p]
        _lib.mpas_b200_launch_count.restype = C.c_longlong
        _lib.mpas_b200_timer_stop.restype = C.c_double
        _lib.mpas_b200_dycore_create.restype = C.c_void_p
        _lib.mpas_b200_dycore_create.argtypes = [C.c_int, C.c_int, C.c_int]
        _lib.mpas_b200_dycore_destroy.argtypes = [C.c_void_p]
        _lib.mpas_b200_dycore_attach_block.argtypes = [C.c_void_p, C.c_int, C.c_void_p]
        _lib.mpas_b200_dycore_set_cfg.argtypes = [C.c_void_p, C.c_int, C.c_char_p, C.c_double]
        _lib.mpas_b200_dycore_get_cfg.restype = C.c_double
        _lib.mpas_b200_dycore_get_cfg.argtypes = [C.c_void_p, C.c_int, C.c_char_p]
        _lib.mpas_b200_nccl_unique_id.argtypes = [C.c_void_p]
        _lib.mpas_b200_dycore_comm_init.argtypes = [C.c_void_p, C.c_void_p]
        _lib.mpas_b200_dycore_init.argtypes = [C.c_void_p]
        _lib.mpas_b200_dycore_step.argtypes = [C.c_void_p, C.c_int]
        _lib.mpas_b200_dycore_get_field.argtypes = [C.c_void_p, C.c_int, C.c_char_p, C.c_void_p, C.c_size_t]
        _lib.mpas_b200_dycore_set_field.argtypes = [C.c_void_p, C.c_int, C.c_char_p, C.c_void_p, C.c_size_t]
        _lib.mpas_b200_dycore_field_device_ptr.restype = C.c_void_p
        _lib.mpas_b200_dycore_field_device_ptr.argtypes = [C.c_void_p, C.c_int, C.c_char_p]
        _lib.mpas_b200_dycore_exchange.argtypes = [C.c_void_p, C.c_char_p, C.c_int]
        _lib.mpas_b200_dycore_call.argtypes = [C.c_void_p, C.c_char_p, C.POINTER(C.c_double), C.c_int]
        _lib.mpas_b200_dycore_exchange_count.restype = C.c_long
        _lib.mpas_b200_dycore_exchange_count.argtypes = [C.c_void_p]
        _lib.mpas_b200_dycore_halo_transport.argtypes = [C.c_void_p]
        _lib.mpas_b200_dycore_set_option.argtypes = [C.c_void_p, C.c_char_p, C.c_int]
        _lib.mpas_b200_dycore_timer_count.argtypes = [C.c_void_p]
        _lib.mpas_b200_dycore_timer_get.argtypes = [C.c_void_p, C.c_int, C.c_char_p, C.c_int, C.POINTER(C.c_double), C.POINTER(C.c_long)]
        _lib.mpas_b200_dycore_timer_reset.argtypes = [C.c_void_p]
        _lib.mpas_b200_dycore_global_max_abs.restype = C.c_double
        _lib.mpas_b200_dycore_global_max_abs.argtypes = [C.c_void_p, C.c_char_p]
        _lib.mpas_b200_dycore_global_minmax.argtypes = [C.c_void_p, C.c_char_p, C.c_longlong, C.c_longlong, C.POINTER(C.c_double), C.POINTER(C.c_double)]
    return _lib


def last_error():
    return lib().mpas_b200_last_error().decode()


def host_error():
    return hostlib().mpas_host_last_error().decode()


class Block:
    """One decomposed block: local mesh + pool of Registry-named fields + exchange lists."""

    def __init__(self, case, handle, iblock):
        self.case = case  # keeps the owning HostCase alive
        self.h = handle
        self.id = iblock
        self._info = None

    def _load_info(self):
        L = hostlib()
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

    def is_lean(self, name):
        """True while `name` is an all-zero work array of a lean block without host storage (field() gives it storage)"""
        return hostlib().mpas_host_block_field_is_lean(self.h, name.encode()) == 1

    def field(self, name):
        kind, loc, dims = self.info(name)
        ptr = hostlib().mpas_host_block_field_ptr(self.h, name.encode())
        if not ptr:
            raise KeyError(name)
        n = int(np.prod(dims))
        ct = C.c_double if kind == 0 else C.c_int32
        arr = np.ctypeslib.as_array(C.cast(ptr, C.POINTER(ct)), shape=(n,))
        return arr.reshape(tuple(reversed(dims)))

    def write_restart(self, path, names):
        """this block's OWNED columns of the named fields into a file made by HostCase.create_restart"""
        if hostlib().mpas_host_block_write_restart(self.h, str(path).encode(), ",".join(names).encode()):
            raise RuntimeError(host_error())

    def read_restart(self, path):
        """every variable of the file that names a pool field (time level 1), owned and halo columns; returns the count"""
        n = hostlib().mpas_host_block_read_restart(self.h, str(path).encode())
        if n < 0:
            raise RuntimeError(host_error())
        return n

    def cfg(self):
        L = hostlib()
        out = {}
        name = C.create_string_buffer(128)
        val = C.c_double()
        for i in range(L.mpas_host_block_ncfg(self.h)):
            L.mpas_host_block_cfg_info(self.h, i, name, 128, C.byref(val))
            out[name.value.decode()] = val.value
        return out

    def exchange_lists(self, kind, which):
        """kind: 'cell'|'edge'|'vertex'; which: 'send'|'recv' -> list of dicts in (endPoint, layer) order."""
        L = hostlib()
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

    def __init__(self, level, lloyd_iters=0, sfc=True, grid_file=None, planar=None):
        """level: icosahedral subdivision level of the built-in mesh generator; grid_file: read an MPAS grid.nc instead;
        planar = (nx, ny, dc): doubly periodic planar mesh of regular hexagons (squall line / supercell: build_block init_case 4 / 5)"""
        if planar is not None:
            self.h = hostlib().mpas_host_case_create_planar(int(planar[0]), int(planar[1]), float(planar[2]), 1 if sfc else 0)
        elif grid_file is not None:
            self.h = hostlib().mpas_host_case_create_from_grid(str(grid_file).encode(), 1 if sfc else 0)
        else:
            self.h = hostlib().mpas_host_case_create(level, lloyd_iters, 1 if sfc else 0)
        if not self.h:
            raise RuntimeError(host_error())
        d = [C.c_int() for _ in range(5)]
        hostlib().mpas_host_case_dims(self.h, *[C.byref(x) for x in d])
        self.nCells, self.nEdges, self.nVertices, self.maxEdges, self.maxEdges2 = [x.value for x in d]
        self.sfc = sfc
        self.nBlocks = 1

    def __del__(self):
        try:
            if self.h:
                hostlib().mpas_host_case_destroy(self.h)
                self.h = None
        except Exception:
            pass

    def write_grid(self, path, version=2):
        """the global mesh as an MPAS grid file (netCDF classic: 1 = CDF-1, 2 = CDF-2 / 64-bit offset, 5 = CDF-5)"""
        if hostlib().mpas_host_case_write_grid(self.h, str(path).encode(), version):
            raise RuntimeError(host_error())

    def create_restart(self, path, names, iblock=0, version=2, xtime="0000-01-01_00:00:00"):
        """header of a restart file for the global mesh; every block then adds its owned columns with Block.write_restart"""
        if hostlib().mpas_host_case_create_restart(self.h, iblock, str(path).encode(), version, ",".join(names).encode(), xtime.encode()):
            raise RuntimeError(host_error())

    def permutation(self, kind):
        n = {"cell": self.nCells, "edge": self.nEdges, "vertex": self.nVertices}[kind]
        out = np.zeros(n, dtype=np.int32)
        if hostlib().mpas_host_case_permutation(self.h, {"cell": 0, "edge": 1, "vertex": 2}[kind], out.ctypes.data):
            raise RuntimeError(host_error())
        return out

    def partition(self, nblocks, method="metis", prefix=""):
        m = {"metis": 0, "sfc": 1, "file": 2}[method]
        if hostlib().mpas_host_case_partition(self.h, nblocks, m, prefix.encode()):
            raise RuntimeError(host_error())
        self.nBlocks = nblocks
        out = np.zeros(self.nCells, dtype=np.int32)
        hostlib().mpas_host_case_get_partition(self.h, out.ctypes.data)
        return out

    def build_block(self, iblock=0, nVertLevels=26, init_case=2, num_scalars=1, full_fields=True, keep_zb=False):
        # full_fields: False = initial state only, True = every dycore field with host storage, "lean" = every dycore field, the
        # all-zero work arrays (83 % of the words) without host storage until somebody asks for them (pool.h Field::zero)
        mode = 2 if full_fields == "lean" else (1 if full_fields else 0)
        rc = hostlib().mpas_host_case_build_block(self.h, iblock, nVertLevels, init_case, num_scalars, mode, 1 if keep_zb else 0)
        if rc:
            raise RuntimeError(host_error())
        return Block(self, hostlib().mpas_host_case_block(self.h, iblock), iblock)


class Dycore:
    """Device-resident dycore (mpas_b200_dycore_*): the atm_srk3 sequence over fields living in HBM.

    blocks: the host-harness Block objects this process owns (one per GPU rank, or several for the
    single-process multi-block mode that mirrors the reference's no-MPI path)."""

    def __init__(self, blocks, cfg=None, rank=0, nranks=1, device=None, diet=True):
        """diet (default): optional physics / output-only arrays get no device mirror and phase-local scratch shares one arena
        per block (include/mpas_b200.h); diet=False gives every pool array its own mirror (routine tests compare them all)."""
        L = lib()
        if L.mpas_b200_device_count() < 1:
            raise RuntimeError("no CUDA device: the dycore has no CPU fallback")
        if device is not None:
            self._chk(L.mpas_b200_set_device(device))
        self.blocks = blocks
        self.h = L.mpas_b200_dycore_create(len(blocks), rank, nranks)
        self.rank, self.nranks = rank, nranks
        self._chk(L.mpas_b200_dycore_set_option(self.h, b"diet", 1 if diet else 0))
        for i, b in enumerate(blocks):
            self._chk(L.mpas_b200_dycore_attach_block(self.h, i, b.h))
            for k, v in (cfg or {}).items():
                L.mpas_b200_dycore_set_cfg(self.h, i, k.encode(), float(v))

    @staticmethod
    def _chk(rc):
        if rc:
            raise RuntimeError("mpas_b200 error %d: %s" % (rc, last_error()))

    def comm_init(self, unique_id_bytes):
        buf = C.create_string_buffer(bytes(unique_id_bytes), 128)
        self._chk(lib().mpas_b200_dycore_comm_init(self.h, buf))

    def init(self):
        self._chk(lib().mpas_b200_dycore_init(self.h))

    def step(self, n=1):
        self._chk(lib().mpas_b200_dycore_step(self.h, n))

    def sync(self):
        self._chk(lib().mpas_b200_synchronize())

    def set_option(self, name, value):
        self._chk(lib().mpas_b200_dycore_set_option(self.h, name.encode(), int(value)))

    def call(self, routine, *params):
        arr = (C.c_double * max(1, len(params)))(*[float(p) for p in params])
        self._chk(lib().mpas_b200_dycore_call(self.h, routine.encode(), arr, len(params)))

    def exchange(self, name, layer=0):
        self._chk(lib().mpas_b200_dycore_exchange(self.h, name.encode(), layer))

    def field(self, ib, name, out=None):
        kind, loc, dims = self.blocks[ib].info(name)
        shape = tuple(reversed(dims))
        if SINGLE and kind == 0:  # device mirror is float32; callers see float64 like the host pools
            tmp = np.empty(shape, dtype=np.float32)
            self._chk(lib().mpas_b200_dycore_get_field(self.h, ib, name.encode(), tmp.ctypes.data, tmp.nbytes))
            if out is None:
                return tmp.astype(np.float64)
            out[...] = tmp
            return out
        if out is None:
            out = np.empty(shape, dtype=np.float64 if kind == 0 else np.int32)
        self._chk(lib().mpas_b200_dycore_get_field(self.h, ib, name.encode(), out.ctypes.data, out.nbytes))
        return out

    def set_field(self, ib, name, arr):
        arr = np.ascontiguousarray(arr)
        if SINGLE and arr.dtype == np.float64:
            arr = arr.astype(np.float32)
        self._chk(lib().mpas_b200_dycore_set_field(self.h, ib, name.encode(), arr.ctypes.data, arr.nbytes))
        self.sync()

    def device_ptr(self, ib, name):
        return lib().mpas_b200_dycore_field_device_ptr(self.h, ib, name.encode())

    def exchange_count(self):
        return lib().mpas_b200_dycore_exchange_count(self.h)

    def halo_transport(self):
        """'none' (one rank), 'nccl' or 'p2p' (stores into the neighbours' mapped receive windows); raises after a p2p timeout"""
        t = lib().mpas_b200_dycore_halo_transport(self.h)
        if t < 0:
            raise RuntimeError("peer-to-peer halo exchange timed out waiting for a neighbour")
        return {0: "none", 1: "nccl", 2: "p2p"}[t]

    def is_scratch_view(self, ib, name):
        """True: `name` is a view of the block's scratch arena (contents valid only inside its phase); None: no device mirror"""
        r = lib().mpas_b200_dycore_field_is_scratch_view(self.h, ib, name.encode())
        return None if r < 0 else bool(r)

    def device_bytes(self):
        return lib().mpas_b200_dycore_device_bytes(self.h)

    def launch_count(self):
        return lib().mpas_b200_launch_count()

    def max_abs(self, name):
        return lib().mpas_b200_dycore_global_max_abs(self.h, name.encode())

    def minmax(self, name, offset=0, stride=1):
        """global (all ranks) min and max over the owned elements, the reductions of summarize_timestep (TI:6675)"""
        lo, hi = C.c_double(), C.c_double()
        self._chk(lib().mpas_b200_dycore_global_minmax(self.h, name.encode(), offset, stride, C.byref(lo), C.byref(hi)))
        return lo.value, hi.value

    def checksum(self, name):
        """sum mod 2^64 of the bit patterns of this process's owned elements (add over ranks: partition independent)"""
        v = C.c_ulonglong()
        self._chk(lib().mpas_b200_dycore_checksum(self.h, name.encode(), C.byref(v)))
        return v.value

    def timers(self):
        L = lib()
        out = {}
        name = C.create_string_buffer(128)
        ms, calls = C.c_double(), C.c_long()
        for i in range(L.mpas_b200_dycore_timer_count(self.h)):
            L.mpas_b200_dycore_timer_get(self.h, i, name, 128, C.byref(ms), C.byref(calls))
            out[name.value.decode()] = (ms.value, calls.value)
        return out

    def reset_timers(self):
        lib().mpas_b200_dycore_timer_reset(self.h)

    def close(self):
        if self.h:
            lib().mpas_b200_dycore_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def nccl_unique_id():
    buf = C.create_string_buffer(128)
    rc = lib().mpas_b200_nccl_unique_id(buf)
    if rc:
        raise RuntimeError(last_error())
    return buf.raw


class NcFile:
    """Classic netCDF (CDF-1/2/5) through the library's own reader / writer (csrc/host/nc_classic.cpp)."""
    TYPES = {1: np.int8, 2: "S1", 3: np.int16, 4: np.int32, 5: np.float32, 6: np.float64}

    def __init__(self, path, mode="r", version=2):
        self.h = hostlib().mpas_nc_open(str(path).encode()) if mode == "r" else hostlib().mpas_nc_create(str(path).encode(), version)
        if not self.h:
            raise RuntimeError(host_error())
        self.mode = mode
        if mode == "r":
            self._inq()

    def _inq(self):
        L = hostlib()
        v, nd, nv, ng = C.c_int(), C.c_int(), C.c_int(), C.c_int()
        nr = C.c_longlong()
        L.mpas_nc_inq(self.h, C.byref(v), C.byref(nd), C.byref(nv), C.byref(ng), C.byref(nr))
        self.version, self.numrecs = v.value, nr.value
        name = C.create_string_buffer(256)
        ln = C.c_longlong()
        self.dims = []
        for i in range(nd.value):
            L.mpas_nc_dim_info(self.h, i, name, 256, C.byref(ln))
            self.dims.append((name.value.decode(), ln.value))
        self.vars = {}
        t, n = C.c_int(), C.c_int()
        ids = (C.c_int * 8)()
        cnt = C.c_longlong()
        for i in range(nv.value):
            L.mpas_nc_var_info(self.h, i, name, 256, C.byref(t), C.byref(n), ids, C.byref(cnt))
            shape = tuple(self.numrecs if self.dims[ids[k]][1] == 0 else self.dims[ids[k]][1] for k in range(n.value))
            self.vars[name.value.decode()] = (t.value, tuple(self.dims[ids[k]][0] for k in range(n.value)), shape, cnt.value)

    def gatt(self, name):
        out = C.create_string_buffer(1024)
        if hostlib().mpas_nc_gatt_text(self.h, name.encode(), out, 1024):
            raise KeyError(name)
        return out.value.decode()

    def get(self, name):
        t, dimnames, shape, cnt = self.vars[name]
        if t in (5, 6):
            out = np.empty(cnt, dtype=np.float64)
            rc = hostlib().mpas_nc_get_double(self.h, name.encode(), out.ctypes.data)
        else:
            out = np.empty(cnt, dtype=np.int32)
            rc = hostlib().mpas_nc_get_int(self.h, name.encode(), out.ctypes.data)
        if rc:
            raise RuntimeError(host_error())
        return out.reshape(shape)

    # ---- writing
    def def_dim(self, name, length):
        i = hostlib().mpas_nc_def_dim(self.h, name.encode(), 0 if length is None else int(length))
        if i < 0:
            raise RuntimeError(host_error())
        return i

    def def_var(self, name, nctype, dimids):
        a = (C.c_int * max(1, len(dimids)))(*dimids)
        i = hostlib().mpas_nc_def_var(self.h, name.encode(), nctype, len(dimids), a)
        if i < 0:
            raise RuntimeError(host_error())
        return i

    def put_gatt(self, name, text):
        if hostlib().mpas_nc_put_gatt_text(self.h, name.encode(), text.encode()):
            raise RuntimeError(host_error())

    def enddef(self, numrecs=0):
        if hostlib().mpas_nc_enddef(self.h, numrecs):
            raise RuntimeError(host_error())

    def put(self, varid, data, rec=0, start=0):
        data = np.ascontiguousarray(data)
        if data.dtype == np.float64:
            rc = hostlib().mpas_nc_put_double(self.h, varid, rec, start, data.size, data.ctypes.data)
        else:
            data = np.ascontiguousarray(data, dtype=np.int32)
            rc = hostlib().mpas_nc_put_int(self.h, varid, rec, start, data.size, data.ctypes.data)
        if rc:
            raise RuntimeError(host_error())

    def close(self):
        if self.h:
            h, self.h = self.h, None
            if hostlib().mpas_nc_close(h):
                raise RuntimeError(host_error())

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass
