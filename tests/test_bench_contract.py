"""bench.py's reference arm runs without a GPU: check the JSON line the driver parses (keys, units, bounded sample)."""
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def run(*extra, env=None):
    out = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--m", "6", "--steps", "1",
                          "--warmup", "1", *extra], capture_output=True, text=True, timeout=300, env=env, cwd=ROOT)
    assert out.returncode == 0, out.stderr[-2000:]
    return [l for l in out.stdout.splitlines() if l.startswith("{")]


def test_reference_arm_line():
    lines = run()
    assert len(lines) == 1
    d = json.loads(lines[0])
    assert d["impl"] == "reference" and d["unit"] == "elements/s" and d["higher_is_better"] is True
    assert d["metric"] == "P2 tet Laplacian form2 assembly throughput"
    assert d["value"] > 0 and d["ms_per_step"] > 0 and d["n_gpus"] == 1 and d["steps"] == 1 and d["warmup"] == 1
    assert d["dtype"] == "f64" and d["vs_baseline"] is None and d["gpu_launches"] == 0
    assert d["config"]["elements"] == 6 * 6 ** 3 and "workload" in d["config"]
    cb = d["cpu_baseline"]
    assert cb["kind"] == "port" and cb["cores"] >= 1 and cb["value"] == d["value"] and "sample" in cb
    assert d["e2e"] == {"value": d["value"], "unit": "elements/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}


def test_reference_arm_under_torchrun_env_prints_on_rank_0_only():
    env = dict(os.environ, RANK="1", WORLD_SIZE="2", LOCAL_RANK="1")
    assert run("--gpus", "2", env=env) == []
    env["RANK"] = "0"
    lines = run("--gpus", "2", env=env)
    assert len(lines) == 1
    d = json.loads(lines[0])
    # the global mesh of the GPU arm at N=2: m grows with N^(1/3)
    assert d["n_gpus"] == 2 and d["config"]["elements"] == 6 * round(6 * 2 ** (1 / 3)) ** 3


def test_reference_arm_on_a_mesh_file(tmp_path):
    """--mesh x.msh: the bench takes a Gmsh file instead of the synthetic cube and reports it as its own workload"""
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import test_gmsh as G
    mesh, tri, marker, ids, perm = G.cube_with_sides(4)
    path = str(tmp_path / "cube4.msh")
    G.write_v2(path, mesh, tri, marker, ids, perm, True)
    lines = run("--mesh", path)
    assert len(lines) == 1
    d = json.loads(lines[0])
    assert d["data"] == "mesh file" and d["config"]["elements"] == 6 * 4 ** 3 and "cube4.msh" in d["config"]["workload"]
    assert "configs[1]" not in d["config"]["workload"] and d["value"] > 0
