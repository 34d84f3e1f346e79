"""Shared test plumbing: load the product host harness, build/load the CPU oracle, and wire a
host case (blocks + pools + exchange lists) into an oracle model."""
import ctypes as C
import importlib.util
import os
import subprocess
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = os.path.join(ROOT, "mpas-model_b200")


def load_pkg_module(name):
    """import mpas-model_b200/<name>.py (the directory name is not a python identifier)."""
    modname = "mpas_model_b200_" + name
    if modname in sys.modules:
        return sys.modules[modname]
    spec = importlib.util.spec_from_file_location(modname, os.path.join(PKG, name + ".py"))
    mod = importlib.util.module_from_spec(spec)
    sys.modules[modname] = mod
    spec.loader.exec_module(mod)
    return mod


host = load_pkg_module("host")

ORACLE_DIR = os.path.join(ROOT, "oracle")
ORACLE_LIB = os.path.join(ORACLE_DIR, "_build", "liboracle.so")
ORACLE_PERF_LIB = os.path.join(ORACLE_DIR, "_build", "liboracle_perf.so")  # -O3 -march=native: the timed CPU baseline
ORACLE_SP_LIB = os.path.join(ORACLE_DIR, "_build", "liboracle_sp.so")  # mpas_real = float (PRECISION=single)
_oracle = None
_oracle_perf = None
_oracle_sp = None

# nominal grid spacing (m) and time step (s) per icosahedral level; config_dt scales with dx from the
# Registry default 720 s at 120 km (Registry.xml nhyd_model record)
LEVEL_DX = {3: 960000.0, 4: 480000.0, 5: 240000.0, 6: 120000.0, 7: 60000.0, 8: 30000.0, 9: 15000.0}


def level_config(level):
    dx = LEVEL_DX[level]
    return dict(config_dt=720.0 * dx / 120000.0, config_len_disp=dx)


def oracle(perf=False, single=False):
    global _oracle, _oracle_perf, _oracle_sp
    if single:
        if _oracle_sp is None:
            if not os.path.exists(ORACLE_SP_LIB):
                subprocess.check_call(["make", "-s", "-C", ORACLE_DIR])
            _oracle_sp = _bind_oracle(C.CDLL(ORACLE_SP_LIB))
            assert _oracle_sp.oracle_real_size() == 4
        return _oracle_sp
    if perf:
        if _oracle_perf is None:
            if not os.path.exists(ORACLE_PERF_LIB):
                subprocess.check_call(["make", "-s", "-C", ORACLE_DIR])
            _oracle_perf = _bind_oracle(C.CDLL(ORACLE_PERF_LIB))
        return _oracle_perf
    if _oracle is None:
        if not os.path.exists(ORACLE_LIB):
            subprocess.check_call(["make", "-s", "-C", ORACLE_DIR])
        _oracle = _bind_oracle(C.CDLL(ORACLE_LIB))
    return _oracle


def _bind_oracle(L):
    if True:
        L.oracle_last_error.restype = C.c_char_p
        L.oracle_model_create.restype = C.c_void_p
        L.oracle_model_create.argtypes = [C.c_int]
        L.oracle_model_destroy.argtypes = [C.c_void_p]
        L.oracle_model_set_field.argtypes = [C.c_void_p, C.c_int, C.c_char_p, C.c_void_p, C.c_int, C.c_int]
        L.oracle_model_set_cfg.argtypes = [C.c_void_p, C.c_int, C.c_char_p, C.c_double]
        L.oracle_model_add_exchange.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p]
        L.oracle_model_exchange.argtypes = [C.c_void_p, C.c_char_p, C.c_int]
        L.oracle_model_init.argtypes = [C.c_void_p]
        L.oracle_model_step.argtypes = [C.c_void_p, C.c_int]
        L.oracle_set_option.argtypes = [C.c_char_p, C.c_int]
        L.oracle_model_call.argtypes = [C.c_void_p, C.c_char_p, C.POINTER(C.c_double), C.c_int]
        L.oracle_model_exchange_count.restype = C.c_long
        L.oracle_model_exchange_count.argtypes = [C.c_void_p]
        L.oracle_model_get_field.restype = C.c_void_p
        L.oracle_model_get_field.argtypes = [C.c_void_p, C.c_int, C.c_char_p]
        L.oracle_model_set_field_f64.argtypes = [C.c_void_p, C.c_int, C.c_char_p, C.c_void_p, C.c_longlong, C.c_int, C.c_int]
        L.oracle_model_timer_count.argtypes = [C.c_void_p]
        L.oracle_model_timer_get.argtypes = [C.c_void_p, C.c_int, C.c_char_p, C.c_int, C.POINTER(C.c_double), C.POINTER(C.c_long)]
        L.oracle_model_timer_reset.argtypes = [C.c_void_p]
        L.oracle_set_num_threads.argtypes = [C.c_int]
    return L


class OracleModel:
    """P virtual ranks of the oracle dycore operating in place on the host blocks' pools."""

    def __init__(self, blocks, cfg, perf=False, single=False):
        """single=True: the float build of the oracle; it keeps its own float copies of the real fields (the host pools stay
        untouched), field() then returns float32 arrays."""
        self.L = oracle(perf, single)
        self.single = single
        self.blocks = blocks
        self.h = self.L.oracle_model_create(len(blocks))
        self._keep = []
        for ib, b in enumerate(blocks):
            for name in b.names():
                kind, loc, dims = b.info(name)
                inner = int(np.prod(dims[:-1])) if len(dims) > 1 else 1
                ptr = host.hostlib().mpas_host_block_field_ptr(b.h, name.encode())
                if kind == 0:
                    self.L.oracle_model_set_field_f64(self.h, ib, name.encode(), ptr, int(np.prod(dims)), inner, loc)
                else:
                    self.L.oracle_model_set_field(self.h, ib, name.encode(), ptr, inner, loc)
            allcfg = dict(b.cfg())
            allcfg.update(cfg)
            for k, v in allcfg.items():
                self.L.oracle_model_set_cfg(self.h, ib, k.encode(), float(v))
            for loc, kind in ((1, "cell"), (2, "edge"), (3, "vertex")):
                for w, which in ((0, "send"), (1, "recv")):
                    for x in b.exchange_lists(kind, which):
                        src = np.ascontiguousarray(x["src"], dtype=np.int32)
                        dst = np.ascontiguousarray(x["dest"], dtype=np.int32)
                        self._keep += [src, dst]
                        self.L.oracle_model_add_exchange(self.h, ib, loc, w, x["endPoint"], x["layer"], len(src), src.ctypes.data, dst.ctypes.data)

    def _chk(self, rc):
        if rc:
            raise RuntimeError(self.L.oracle_last_error().decode())

    def init(self):
        self._chk(self.L.oracle_model_init(self.h))

    def step(self, n=1):
        self._chk(self.L.oracle_model_step(self.h, n))

    def exchange(self, name, layer=0):
        self._chk(self.L.oracle_model_exchange(self.h, name.encode(), layer))

    def call(self, routine, *params):
        arr = (C.c_double * max(1, len(params)))(*[float(p) for p in params])
        self._chk(self.L.oracle_model_call(self.h, routine.encode(), arr, len(params)))

    def field(self, ib, name):
        """numpy view of the CURRENT array behind `name` (time levels swap every step)."""
        b = self.blocks[ib]
        kind, loc, dims = b.info(name)
        ptr = self.L.oracle_model_get_field(self.h, ib, name.encode())
        n = int(np.prod(dims))
        ct = (C.c_float if self.single else C.c_double) if kind == 0 else C.c_int32
        return np.ctypeslib.as_array(C.cast(ptr, C.POINTER(ct)), shape=(n,)).reshape(tuple(reversed(dims)))

    def exchange_count(self):
        return self.L.oracle_model_exchange_count(self.h)

    def timers(self):
        """{reference timer name: (ms, calls)} of the steps taken since reset_timers (mpas_timer names, TI:341...1369)"""
        out = {}
        name = C.create_string_buffer(128)
        ms, calls = C.c_double(), C.c_long()
        for i in range(self.L.oracle_model_timer_count(self.h)):
            self.L.oracle_model_timer_get(self.h, i, name, 128, C.byref(ms), C.byref(calls))
            out[name.value.decode()] = (ms.value, calls.value)
        return out

    def reset_timers(self):
        self.L.oracle_model_timer_reset(self.h)

    def __del__(self):
        try:
            self.L.oracle_model_destroy(self.h)
        except Exception:
            pass


class GpuModel(host.Dycore):
    """The product dycore on cuda:0 (or the given device) for the given host blocks."""

    def __init__(self, blocks, cfg, **kw):
        super().__init__(blocks, dict(cfg), **kw)  # per-block dims/configs come from attach_block


def has_gpu():
    try:
        return host.lib().mpas_b200_device_count() > 0
    except Exception:
        return False


def oracle_set_option(name, value):
    """cr_pow=1: the oracle evaluates exner with a correctly rounded pow instead of libm's (see oracle/dycore_oracle.cpp)"""
    assert oracle().oracle_set_option(name.encode(), int(value)) == 0


def make_case(level, nblocks=1, nVertLevels=26, init_case=2, num_scalars=1, method="sfc", lloyd=10, sfc=True):
    case = host.HostCase(level, lloyd_iters=lloyd, sfc=sfc)
    if nblocks > 1:
        case.partition(nblocks, method)
    blocks = [case.build_block(i, nVertLevels=nVertLevels, init_case=init_case, num_scalars=num_scalars, full_fields=True) for i in range(nblocks)]
    return case, blocks


def rel_l2(a, b):
    a = np.asarray(a, dtype=np.float64).ravel()
    b = np.asarray(b, dtype=np.float64).ravel()
    den = np.sqrt((b * b).sum())
    num = np.sqrt(((a - b) ** 2).sum())
    return num / den if den > 0 else num
