"""Host-compiled unit tests of the device headers (no GPU):
  * csrc/fft_exact.cuh emulated thread by thread == the oracle's kissfft restatement, SF 2..12,
    and the shared-memory exchange pattern is bank-conflict free;
  * csrc/glibc_math.cuh == the running glibc over a strided walk of all float bit patterns."""
import json
import os
import subprocess

import pytest

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
CSRC = os.path.join(ROOT, "lora-sdr-lightweight-standalone-library-clean_b200", "csrc")
BUILD = os.path.join(HERE, "cpu", "build")


def compile_cpp(name, extra):
    os.makedirs(BUILD, exist_ok=True)
    out = os.path.join(BUILD, name)
    src = os.path.join(HERE, "cpu", name + ".cpp")
    cmd = ["g++", "-std=c++17", "-O2", "-ffp-contract=off", "-pthread", "-I", CSRC, "-o", out, src] + extra
    subprocess.run(cmd, check=True)
    return out


def test_fft_emulation_matches_oracle(oracle):
    exe = compile_cpp("test_fft_emulation", ["-x", "c", os.path.join(ROOT, "oracle", "lora_oracle.c"), "-x", "none", "-lm"])
    r = subprocess.run([exe], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    lines = [json.loads(l) for l in r.stdout.splitlines() if l.startswith("{")]
    assert len(lines) >= 11
    for l in lines:
        assert l.get("match", True)


def test_glibc_math_restatement():
    exe = compile_cpp("test_glibc_math", ["-lm"])
    r = subprocess.run([exe, "1031"], capture_output=True, text=True, timeout=600)   # ~4.2 M arguments per function
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]


def test_fft_fast_emulation():
    """csrc/fft_fast.cuh and fft_soa.cuh (the certified path's throughput transform) emulated thread by thread: spectrum
    within 2e-6 of a double DFT, every exchange address written once and read once inside the window slot, every 64-bit
    access phase bank-conflict free, output bin map a permutation -- SF7 (both register layouts) to SF12; and the pruned
    transform of fft_prune.cuh (pass-0 and geometric variants, single- and multi-warp windows): every bin of every class
    against the double DFT, bin map covering each class exactly once; and its streaming form (demod_lean.cu: R accumulators
    per lane, chunk joins from the table's powers of two, in-register cross stage) for SF10-12."""
    exe = compile_cpp("test_fft_fast_emulation", ["-lm"])
    r = subprocess.run([exe], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    lines = [json.loads(l) for l in r.stdout.splitlines() if l.startswith("{")]
    assert len(lines) == 18 and all(l["match"] and l["worst_bank_conflict"] == 1 for l in lines)
    assert sum("pruned" in l.get("layout", "") for l in lines) == 11
