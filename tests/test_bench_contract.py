"""The reference arm of bench.py runs on host cores only (the oracle port of week-2/diet/simplex.py): its JSON line
can be checked here without a GPU.  Contract: one line, the same metric / unit / config as the GPU arm,
"impl": "reference", a cpu_baseline describing the run and an e2e block with zero copy bytes; ranks other than 0
print nothing and exit 0."""
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _run(extra_env=None):
    env = dict(os.environ)
    env.pop("RANK", None)
    env.pop("WORLD_SIZE", None)
    env.update(extra_env or {})
    return subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--steps", "1",
                           "--warmup", "0"], capture_output=True, text=True, env=env, timeout=600)


def test_reference_arm_prints_one_contract_line():
    out = _run()
    assert out.returncode == 0, out.stderr[-2000:]
    lines = [ln for ln in out.stdout.splitlines() if ln.startswith("{")]
    assert len(lines) == 1
    d = json.loads(lines[0])
    assert d["impl"] == "reference" and d["metric"].startswith("LPs/sec") and d["unit"] == "LPs/s"
    assert d["higher_is_better"] is True and d["value"] > 0 and d["steps"] == 1 and d["n_gpus"] == 1
    assert d["config"]["workload"].startswith("configs[3]") and d["config"]["n"] == 100 and d["config"]["m"] == 100
    assert d["cpu_baseline"]["kind"] == "port" and d["cpu_baseline"]["cores"] >= 1 and d["cpu_baseline"]["value"] == d["value"]
    assert d["e2e"] == {"value": d["value"], "unit": "LPs/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}
    assert d["gpu_launches"] == 0 and d["vs_baseline"] is None and d["dtype"] == "f64"


def test_reference_arm_other_ranks_stay_silent():
    out = _run({"RANK": "1", "WORLD_SIZE": "2", "LOCAL_RANK": "1"})
    assert out.returncode == 0, out.stderr[-2000:]
    assert not [ln for ln in out.stdout.splitlines() if ln.startswith("{")]


def test_ncu_facts_are_reported_only_for_the_build_they_were_taken_from(tmp_path, monkeypatch):
    """bench.py quotes ncu numbers (`traffic`, wavefronts per pivot, DRAM bytes per system) from profiles/ncu_facts.json
    only while the hash recorded with the capture equals the hash of the kernel's own source files NOW; a capture of
    another build is reported as stale (-> `traffic: null`), never silently."""
    import importlib
    import json
    sys.path.insert(0, ROOT)
    bench = importlib.import_module("bench")
    for k in ("simplex_batched_kernel", "simplex_single_kernel", "gauss_jordan_batched_kernel"):
        h = bench.source_sha16(k)
        assert len(h) == 16 and h != bench.source_sha16()          # per-kernel hash differs from the whole-tree one
        assert h == bench.source_sha16(k)                            # deterministic
    fake_root = tmp_path
    (fake_root / "profiles").mkdir()
    csrc = fake_root / "linear_programming_b200" / "csrc"
    csrc.mkdir(parents=True)
    for name in set(sum(bench.KERNEL_SOURCES.values(), [])):
        (csrc / name).write_text("// " + name + "\n")
    monkeypatch.setattr(bench, "ROOT", str(fake_root))
    fresh = bench.source_sha16("simplex_single_kernel")
    facts = {"simplex_single_kernel": {"src_sha16": fresh, "dram_bytes_per_pivot": 8.59e9, "file": "x.csv"},
             "gauss_jordan_batched_kernel": {"src_sha16": "0123456789abcdef", "dram_bytes_per_system": 1.0, "file": "y.csv"}}
    (fake_root / "profiles" / "ncu_facts.json").write_text(json.dumps(facts))
    assert bench.ncu_facts("simplex_single_kernel")["dram_bytes_per_pivot"] == 8.59e9
    assert bench.ncu_facts("gauss_jordan_batched_kernel")["stale"] is True
    assert bench.ncu_facts("simplex_batched_kernel") is None          # no capture at all
    (csrc / "simplex_single.cu").write_text("// edited\n")             # the kernel changes -> its facts go stale
    assert bench.ncu_facts("simplex_single_kernel")["stale"] is True
    (csrc / "host_narrow.cpp").write_text("// host-side edit\n")       # ... but not another file's edit
    assert bench.ncu_facts("gauss_jordan_batched_kernel")["captured_src_sha16"] == "0123456789abcdef"
