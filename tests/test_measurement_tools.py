"""The scripts that turn ncu captures into the numbers under profiles/ (CPU only, synthetic captures)."""
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

HDR = '"ID","Process ID","Process Name","Host Name","Kernel Name","Context","Stream","Block Size","Grid Size","Device","CC","Section Name","Metric Name","Metric Unit","Metric Value"'


def _row(i, kernel, metric, unit, value):
    return '"%d","1","python","box","%s","1","7","(256, 1, 1)","(10, 1, 1)","0","10.0","Command line profiler metrics","%s","%s","%s"' % (i, kernel, metric, unit, value)


def _capture(path, launches):
    """launches: [(kernel, us, read MB, written MB)] -> a CSV with three metrics per launch, as the traffic pass of ncu writes it"""
    lines = ["==PROF== Connected to process 1", HDR]
    for i, (k, us, rd, wr) in enumerate(launches):
        lines.append(_row(i, k, "dram__bytes_read.sum", "Mbyte", "%.3f" % rd))
        lines.append(_row(i, k, "dram__bytes_write.sum", "Mbyte", "%.3f" % wr))
        lines.append(_row(i, k, "gpu__time_duration.sum", "us", "%.3f" % us))
    open(path, "w").write("\n".join(lines) + "\n")


def test_launch_summary_uses_only_the_time_metric_of_a_multi_metric_capture(tmp_path):
    """a capture with dram bytes beside the time of every launch: launches are counted once and bytes never enter the time sums
    (the round-2 summary once counted every launch three times and summed megabytes into microseconds)"""
    cap = tmp_path / "cap.csv"
    _capture(str(cap), [("void <unnamed>::a_kernel<1>(P)", 100.0, 1000.0, 500.0)] * 3 + [("<unnamed>::b_kernel(P)", 50.0, 10.0, 10.0)] * 2)
    out = subprocess.run([sys.executable, os.path.join(ROOT, "tools", "summarize_launches.py"), str(cap)], capture_output=True, text=True, check=True).stdout
    rows = [l.split(",") for l in out.strip().splitlines()]
    a = next(r for r in rows if "a_kernel" in r[0])
    b = next(r for r in rows if "b_kernel" in r[0])
    tot = rows[-1]
    assert a[1:4] == ["3", "300.0", "100.0"] and b[1:4] == ["2", "100.0", "50.0"]
    assert tot[0] == "TOTAL" and tot[1] == "5" and tot[2] == "400.0"
    assert abs(float(a[4]) - 75.0) < 0.01


def test_dram_traffic_json_carries_the_source_hash_and_the_workload(tmp_path):
    """tools/dram_traffic.py stamps its json with the hash of the kernel sources and the workload; bench.py uses the capture for
    roofline.traffic only when the hash equals the sources it runs (bench.py: kernel_source_sha)"""
    cap = tmp_path / "cap.csv"
    _capture(str(cap), [("void <unnamed>::dt_edge_kernel<5, 0, 0, 5>(<unnamed>::DtP)", 600.0, 1180.0, 205.0)] * 7 + [("<unnamed>::recover_cell_kernel(P)", 245.0, 760.0, 530.0)] * 7)
    js = tmp_path / "t.json"
    subprocess.run([sys.executable, os.path.join(ROOT, "tools", "dram_traffic.py"), str(cap), str(js), "7", "55", "4"], capture_output=True, text=True, check=True)
    d = json.load(open(js))
    sys.path.insert(0, ROOT)
    import bench
    assert d["_kernel_source_sha"] == bench.kernel_source_sha()
    assert (d["_nCells"], d["_nVertLevels"], d["_num_scalars"]) == (163842, 55, 4)
    # 7 steps in the capture (tools/dram_traffic.py NSTEPS): one launch per step here
    assert abs(d["atm_compute_dyn_tend"]["dram_GB_per_step"] - 1.385) < 1e-6
    assert abs(d["atm_recover_large_step_variables"]["dram_GB_per_step"] - 1.290) < 1e-6
