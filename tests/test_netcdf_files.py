"""On-disk formats: the library's own classic-netCDF reader / writer against an independent
implementation (scipy.io.netcdf_file, CDF-1 and CDF-2), CDF-5 round trip, the MPAS grid file round trip, and the
restart idiom of the reference: run, write a restart, read it into a fresh case, continue -> bit-identical to the
uninterrupted run."""
import numpy as np
import pytest
from scipy.io import netcdf_file

import common

host = common.host


def test_writer_is_read_by_scipy(tmp_path):
    rng = np.random.default_rng(3)
    a = rng.standard_normal((5, 7))
    b = rng.integers(-2 ** 31, 2 ** 31 - 1, size=11, dtype=np.int64).astype(np.int32)
    r0, r1 = rng.standard_normal((2, 3, 4))
    for version in (1, 2):
        p = tmp_path / ("w%d.nc" % version)
        f = host.NcFile(p, "w", version)
        dT = f.def_dim("Time", None)
        dx, dy, dz = f.def_dim("x", 5), f.def_dim("y", 7), f.def_dim("z", 11)
        d3, d4 = f.def_dim("p", 3), f.def_dim("q", 4)
        va = f.def_var("a", 6, [dx, dy])
        vb = f.def_var("b", 4, [dz])
        vr = f.def_var("r", 6, [dT, d3, d4])
        vs = f.def_var("s", 4, [dT])
        f.put_gatt("title", "classic netcdf test")
        f.enddef(2)
        f.put(va, a)
        f.put(vb, b)
        f.put(vr, r0, rec=0)
        f.put(vr, r1, rec=1)
        f.put(vs, np.array([7], dtype=np.int32), rec=0)
        f.put(vs, np.array([-9], dtype=np.int32), rec=1)
        f.close()
        g = netcdf_file(str(p), "r", mmap=False)
        assert g.version_byte == version
        assert g.dimensions["x"] == 5 and g.dimensions["Time"] is None
        assert g.title == b"classic netcdf test"
        assert np.array_equal(g.variables["a"][:], a)
        assert np.array_equal(g.variables["b"][:], b)
        assert np.array_equal(g.variables["r"][:], np.stack([r0, r1]))
        assert np.array_equal(g.variables["s"][:], [7, -9])
        g.close()


def test_reader_reads_scipy_files(tmp_path):
    rng = np.random.default_rng(5)
    a = rng.standard_normal((4, 6))
    i = np.arange(-3, 6, dtype=np.int32)
    f32 = rng.standard_normal(5).astype(np.float32)
    rec = rng.standard_normal((3, 2, 5))
    for version in (1, 2):
        p = tmp_path / ("s%d.nc" % version)
        g = netcdf_file(str(p), "w", version=version)
        g.history = "made by scipy"
        g.createDimension("Time", None)
        g.createDimension("x", 4); g.createDimension("y", 6); g.createDimension("n", 9); g.createDimension("m", 5); g.createDimension("two", 2)
        g.createVariable("a", "d", ("x", "y"))[:] = a
        g.createVariable("i", "i", ("n",))[:] = i
        g.createVariable("f", "f", ("m",))[:] = f32
        v = g.createVariable("rec", "d", ("Time", "two", "m"))
        v[0:3] = rec
        w = g.createVariable("t", "i", ("Time",))
        w[0:3] = [1, 2, 3]
        g.close()
        f = host.NcFile(p)
        assert f.version == version and f.numrecs == 3
        assert dict(f.dims)["Time"] == 0 and dict(f.dims)["y"] == 6
        assert f.gatt("history") == "made by scipy"
        assert np.array_equal(f.get("a"), a)
        assert np.array_equal(f.get("i"), i)
        assert np.array_equal(f.get("f"), f32.astype(np.float64))
        assert np.array_equal(f.get("rec"), rec)
        assert np.array_equal(f.get("t"), [1, 2, 3])
        f.close()


def test_cdf5_round_trip_and_rejects_other_files(tmp_path):
    a = np.linspace(-1, 1, 35).reshape(5, 7)
    p = tmp_path / "five.nc"
    f = host.NcFile(p, "w", 5)
    v = f.def_var("a", 6, [f.def_dim("x", 5), f.def_dim("y", 7)])
    f.enddef()
    f.put(v, a)
    f.close()
    assert open(p, "rb").read(4) == b"CDF\x05"
    g = host.NcFile(p)
    assert g.version == 5 and np.array_equal(g.get("a"), a)
    g.close()
    bad = tmp_path / "bad.nc"
    bad.write_bytes(b"\x89HDF\r\n\x1a\n" + b"\0" * 64)
    with pytest.raises(RuntimeError, match="not a netCDF classic file"):
        host.NcFile(bad)


def test_grid_file_round_trip_gives_identical_blocks(tmp_path):
    case = host.HostCase(3, lloyd_iters=4, sfc=True)
    p = tmp_path / "x1.642.grid.nc"
    case.write_grid(p)
    g = netcdf_file(str(p), "r", mmap=False)  # an independent reader sees an MPAS-style grid file
    assert g.dimensions["nCells"] == 642 and g.dimensions["nEdges"] == 1920 and g.dimensions["vertexDegree"] == 3
    assert g.on_a_sphere == b"YES" and g.variables["cellsOnEdge"].shape == (1920, 2)
    assert g.variables["cellsOnEdge"][:].min() >= 1 and g.variables["cellsOnEdge"][:].max() <= 642
    g.close()
    case2 = host.HostCase(0, grid_file=p, sfc=False)  # the file already is in SFC order
    assert (case2.nCells, case2.nEdges, case2.nVertices) == (642, 1920, 1280)
    for c in (case, case2):
        c.partition(2, "sfc")
    for ib in range(2):
        b1 = case.build_block(ib, nVertLevels=8, num_scalars=2)
        b2 = case2.build_block(ib, nVertLevels=8, num_scalars=2)
        assert b1.names() == b2.names()
        for n in b1.names():
            assert np.array_equal(b1.field(n), b2.field(n)), n


RESTART = ["u@1", "w@1", "theta_m@1", "rho_zz@1", "scalars@1", "ru", "rw", "rho_p", "rtheta_p", "exner", "pressure_p", "rho_edge", "pv_edge", "v", "ke",
           "divergence", "vorticity",
           "rtheta_base", "exner_base", "pressure_base"]  # base state that atm_init_coupled_diagnostics derives (skipped on restart, mpas_atm_core.F:389)


def _tracers(blocks):
    for b in blocks:
        s = b.field("scalars@1")
        lat = b.field("latCell")
        s[:-1, :, 1] = (np.sin(2 * lat) ** 2)[:-1, None]


def test_restart_run_is_bit_identical_to_the_uninterrupted_run(tmp_path):
    cfg = common.level_config(3)
    case, blocks = common.make_case(3, 2, nVertLevels=8, num_scalars=2)
    _tracers(blocks)
    ref = common.OracleModel(blocks, cfg)
    ref.init()
    ref.step(2)
    p = tmp_path / "restart.nc"
    case.create_restart(p, RESTART, xtime="0000-01-01_01:00:00")
    for b in blocks:
        b.write_restart(p, RESTART)
    saved = [{n: ref.field(ib, n).copy() for n in RESTART} for ib in range(2)]
    ref.step(2)

    g = netcdf_file(str(p), "r", mmap=False)  # global order, Registry-style names and dimensions
    assert g.variables["u"].dimensions == ("Time", "nEdges", "nVertLevels")
    assert g.variables["w"].dimensions == ("Time", "nCells", "nVertLevelsP1")
    assert g.variables["qv"].shape == (1, 642, 8) and "tracer2" in g.variables
    assert bytes(g.variables["xtime"][0]).decode().strip() == "0000-01-01_01:00:00"
    for ib, b in enumerate(blocks):
        nown = int(b.cfg()["nEdgesSolve"])
        ids = b.field("indexToEdgeID")[:nown] - 1
        assert np.array_equal(g.variables["u"][0][ids], saved[ib]["u@1"][:nown])
    g.close()

    case2, blocks2 = common.make_case(3, 2, nVertLevels=8, num_scalars=2)
    SOLVE = {1: "nCellsSolve", 2: "nEdgesSolve", 3: "nVerticesSolve"}
    IDS = {1: "indexToCellID", 2: "indexToEdgeID", 3: "indexToVertexID"}
    for ib, b in enumerate(blocks2):
        assert b.read_restart(p) == len(RESTART)
    for ib, b in enumerate(blocks2):
        for n in RESTART:
            loc = b.info(n)[1]
            nown = int(b.cfg()[SOLVE[loc]])
            assert np.array_equal(b.field(n)[:nown], saved[ib][n][:nown]), n
            # halo columns hold their owners' values (the run itself leaves e.g. the scalar halo stale at a step end)
            owner = {int(g): (jb, j) for jb, bb in enumerate(blocks2) for j, g in enumerate(bb.field(IDS[loc])[:int(bb.cfg()[SOLVE[loc]])])}
            ids = b.field(IDS[loc])[:-1]
            for j in (nown, len(ids) - 1):
                jb, jo = owner[int(ids[j])]
                assert np.array_equal(b.field(n)[j], saved[jb][n][jo]), (n, j)
    cfg2 = dict(cfg)
    cfg2["config_do_restart"] = 1
    m = common.OracleModel(blocks2, cfg2)
    m.init()
    m.step(2)
    for ib, b in enumerate(blocks2):
        for n in RESTART:
            nown = int(b.cfg()[SOLVE[b.info(n)[1]]])
            assert np.array_equal(m.field(ib, n)[:nown], ref.field(ib, n)[:nown]), n


@pytest.mark.gpu
def test_gpu_restart_run_is_bit_identical_to_the_uninterrupted_run(tmp_path):
    cfg = common.level_config(4)
    case, blocks = common.make_case(4, 2, nVertLevels=26, num_scalars=2)
    _tracers(blocks)
    ref = common.GpuModel(blocks, cfg)
    ref.init()
    ref.step(3)  # an odd number of steps: the device has swapped its time levels, the download below follows them
    for ib, b in enumerate(blocks):
        for n in RESTART:
            b.field(n)[...] = ref.field(ib, n)
    p = tmp_path / "restart_gpu.nc"
    case.create_restart(p, RESTART)
    for b in blocks:
        b.write_restart(p, RESTART)
    ref.step(2)
    case2, blocks2 = common.make_case(4, 2, nVertLevels=26, num_scalars=2)
    for b in blocks2:
        assert b.read_restart(p) == len(RESTART)
    cfg2 = dict(cfg)
    cfg2["config_do_restart"] = 1
    m = common.GpuModel(blocks2, cfg2)
    m.init()
    m.step(2)
    SOLVE = {1: "nCellsSolve", 2: "nEdgesSolve", 3: "nVerticesSolve"}
    for ib, b in enumerate(blocks2):
        for n in RESTART:
            nown = int(b.cfg()[SOLVE[b.info(n)[1]]])
            assert np.array_equal(m.field(ib, n)[:nown], ref.field(ib, n)[:nown]), n


def test_restart_of_another_mesh_or_level_count_is_refused(tmp_path):
    """variables are matched by name: a file whose sizes merely divide evenly must not be read from the wrong offsets"""
    case3, blocks3 = common.make_case(3, 1, nVertLevels=8, num_scalars=1)
    p = tmp_path / "restart_l3.nc"
    case3.create_restart(p, ["u@1", "theta_m@1", "w@1"])
    blocks3[0].write_restart(p, ["u@1", "theta_m@1", "w@1"])
    assert blocks3[0].read_restart(p) == 3
    case4, blocks4 = common.make_case(4, 1, nVertLevels=8, num_scalars=1)
    with pytest.raises(RuntimeError, match="another mesh|too small"):
        blocks4[0].read_restart(p)
    case3b, blocks3b = common.make_case(3, 1, nVertLevels=4, num_scalars=1)
    with pytest.raises(RuntimeError, match="nVertLevels"):
        blocks3b[0].read_restart(p)
