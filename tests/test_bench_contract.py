"""bench.py's reference arm and the pieces the driver pairs the two arms by (CPU only; the GPU arm runs on the box)."""
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402


def _ref_line(workload, steps=2, env_extra=None):
    env = dict(os.environ, **(env_extra or {}))
    r = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--workload", workload,
                        "--steps", str(steps), "--warmup", "3"], capture_output=True, text=True, timeout=300, env=env, cwd=ROOT)
    assert r.returncode == 0, r.stderr[-2000:]
    lines = [ln for ln in r.stdout.splitlines() if ln.startswith("{")]
    assert len(lines) == 1
    return json.loads(lines[0])


def test_reference_arm_line_pairs_with_the_repo_arm():
    """Same metric string and config as the repo arm would print, real steps (steps x ms_per_step is the time actually
    spent), all host cores even when the launcher exported OMP_NUM_THREADS=1 (torch.distributed.run does)."""
    w = bench.WORKLOADS["burgers3d_256"]
    d = _ref_line("burgers3d_256", steps=2, env_extra={"OMP_NUM_THREADS": "1"})
    assert d["impl"] == "reference" and d["metric"] == bench.metric_name(w, "f64") and d["unit"] == "Mcell-updates/s"
    assert d["config"] == bench.make_config("burgers3d_256", w, 1, "f64", "peer")
    assert d["higher_is_better"] is True and d["steps"] == 2 and d["warmup"] == 3
    cb = d["cpu_baseline"]
    assert cb["kind"] == "port" and cb["value"] == d["value"] == d["e2e"]["value"]
    assert cb["cores"] == len(os.sched_getaffinity(0))
    assert abs(d["ms_per_step"] * d["steps"] * 1e-3 - cb["seconds"]) < 1e-6
    assert "matlab_port.c" in cb["sample"] and "cuda_program_restatement" in cb
    assert d["e2e"]["h2d_bytes_per_step"] == 0 and d["e2e"]["d2h_bytes_per_step"] == 0


def test_reference_arm_other_ranks_do_nothing():
    env = dict(os.environ, RANK="1", WORLD_SIZE="2", LOCAL_RANK="1")
    r = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--gpus", "2", "--steps", "2"],
                       capture_output=True, text=True, timeout=120, env=env, cwd=ROOT)
    assert r.returncode == 0 and r.stdout.strip() == ""


def test_capture_hash_is_keyed_on_the_captured_instantiations():
    sys.path.insert(0, os.path.join(ROOT, "tools"))
    from sass_hashes import family_hash
    built = {"_Z1kIdLi1ELb0EE": "aa", "_Z1kIdLi2ELb0EE": "bb", "_Z1kIdLi2ELb1EE": "cc"}
    h = family_hash(built, r"_Z1kIdLi[12]ELb0E")
    assert h == family_hash(dict(built, _Z1kIfLi1ELb0EE="dd"), r"_Z1kIdLi[12]ELb0E")        # another instantiation: no effect
    assert h != family_hash(dict(built, _Z1kIdLi2ELb0EE="xx"), r"_Z1kIdLi[12]ELb0E")        # a captured one changed
    assert family_hash(built, r"nothing") is None


def test_every_workload_names_a_known_target():
    import multigpu_advectiondiffusion_b200 as adv
    for name, w in bench.WORKLOADS.items():
        assert w["target"] in adv.TARGETS, name
        assert bench.metric_name(w, "f64").startswith("Mcell-updates/s (")
