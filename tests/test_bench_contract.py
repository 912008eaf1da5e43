"""bench.py contract pieces that can be checked without a GPU."""
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_reference_arm_reports_unavailable_with_one_json_line():
    r = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--gpus", "1", "--steps", "2",
                        "--warmup", "1"], capture_output=True, text=True, timeout=120, cwd=ROOT)
    assert r.returncode == 0
    lines = [l for l in r.stdout.splitlines() if l.strip()]
    assert len(lines) == 1
    d = json.loads(lines[0])
    assert d["impl"] == "reference" and "unavailable" in d and "tensorflow" in d["unavailable"].lower()


def test_graft_entry_build_produces_the_in_tree_library():
    sys.path.insert(0, ROOT)
    import __graft_entry__
    __graft_entry__.build()
    from tensorflow_distributed_b200.ops import native
    assert os.path.exists(native.lib_path()) and native.lib_path().startswith(ROOT)
    assert native.lib().tfd_version() == 100


def test_performance_table_schema_matches_the_reference_file():
    from tensorflow_distributed_b200.utils.timing import performance_table
    txt = performance_table([(40, 200.0, 90.0, 0.01), (60, 290.0, 93.0, 0.01)])
    assert txt.splitlines()[0] == "Steps ,Time ,Accuracy, Learning rate"  # /root/reference/performance:1
    assert txt.splitlines()[1].split()[0] == "40" and txt.splitlines()[2].split()[-1] == "0.01"


def test_committed_performance_record_has_the_reference_schema():
    """`performance` at the repo root mirrors /root/reference/performance:1-6: a header and five rows of
    steps / time / accuracy / learning rate at 40..120 steps."""
    import os
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    lines = [l for l in open(os.path.join(root, "performance")).read().splitlines() if l and not l.startswith("#")]
    assert lines[0] == "Steps ,Time ,Accuracy, Learning rate"
    rows = [l.split() for l in lines[1:]]
    assert [int(r[0]) for r in rows] == [40, 60, 80, 100, 120]
    times = [float(r[1]) for r in rows]
    assert times == sorted(times) and all(float(r[3]) == 0.01 for r in rows)
    assert all(0.0 <= float(r[2]) <= 100.0 for r in rows)
