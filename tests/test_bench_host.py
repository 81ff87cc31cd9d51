"""Host-side pieces of bench.py that run without a GPU: the reference arm's JSON line (the contract the driver
parses) and the best-effort CPU binding of multi-rank runs."""
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_bind_near_gpu_never_fails_and_never_widens_the_affinity():
    sys.path.insert(0, ROOT)
    import bench
    before = os.sched_getaffinity(0)
    try:
        got = bench.bind_near_gpu(0)            # no GPU / no NVML here: None; on a GPU box: a subset of `before`
        after = os.sched_getaffinity(0)
        assert after <= before and len(after) > 0
        assert got is None or (got["cpus"] == len(after) and got["first"] in after and got["last"] in after)
    finally:
        os.sched_setaffinity(0, before)


def test_reference_arm_prints_one_contract_line():
    """`bench.py --impl reference` runs the compiled reference (oracle/_ref) on host threads only: same metric / unit /
    config keys as the GPU arm, `impl: reference`, an e2e block with zero copy bytes, no GPU needed."""
    if not os.path.exists(os.path.join(ROOT, "oracle", "_ref", "libref_lora_phy.so")):
        import pytest
        pytest.skip("oracle/_ref not built (python -c 'import __graft_entry__ as g; g.build()')")
    # 20 steps of the smallest sample the arm sizes (2 s of work each by its own rule)
    out = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--steps", "20", "--warmup", "0"],
                         capture_output=True, text=True, timeout=600, cwd=ROOT)
    assert out.returncode == 0, out.stderr[-2000:]
    lines = [l for l in out.stdout.splitlines() if l.startswith("{")]
    assert len(lines) == 1
    rec = json.loads(lines[0])
    assert rec["impl"] == "reference" and rec["metric"] == "demod symbols/s" and rec["unit"] == "symbols/s"
    assert rec["higher_is_better"] is True and rec["value"] > 0
    assert rec["e2e"]["h2d_bytes_per_step"] == 0 and rec["e2e"]["d2h_bytes_per_step"] == 0
    assert rec["cpu_baseline"]["kind"] in ("reference", "port") and rec["cpu_baseline"]["cores"] >= 1
    assert "workload" in rec["config"]
