"""tools/fuzz_host.cpp under AddressSanitizer + UBSan for a few seconds: the host-side parsers (tuple decoder, message
encoder, Float.toString printer, SavedModel importer, blob parser) take mutated input without touching a byte outside
their buffers. A longer run's result is kept in profiles/fuzz_host_r02.json."""
import json
import os
import shutil
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CSRC = os.path.join(ROOT, "distributed-inference-system-based-storm_b200", "csrc")


@pytest.mark.timeout(300)
def test_host_parsers_survive_mutated_input(tmp_path):
    gxx = shutil.which("g++")
    if not gxx:
        pytest.skip("no g++")
    exe = str(tmp_path / "fuzz_host")
    cmd = [gxx, "-std=c++17", "-O1", "-g", "-fsanitize=address,undefined", "-fno-sanitize-recover=undefined",
           "-I", os.path.join(ROOT, "include"), os.path.join(ROOT, "tools", "fuzz_host.cpp"),
           *[os.path.join(CSRC, f) for f in ("sib_json.cpp", "sib_pbimport.cpp", "sib_model.cpp")], "-o", exe]
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode and ("asan" in r.stderr.lower() or "sanitize" in r.stderr.lower()):
        pytest.skip("this g++ has no sanitizer runtime")
    assert r.returncode == 0, r.stderr[-2000:]
    seed = os.path.join(ROOT, "tests", "golden", "lenet_tf113_saved_model.pb")
    r = subprocess.run([exe, "5", seed], capture_output=True, text=True, env={**os.environ, "ASAN_OPTIONS": "detect_leaks=1"})
    assert r.returncode == 0, (r.stdout[-500:], r.stderr[-3000:])
    out = json.loads(r.stdout.strip().splitlines()[-1])
    assert out["findings"] == 0 and out["rounds"] > 1000
    assert out["json_decode_ok"] > 0 and out["json_decode_rejected"] > 0 and out["import_pb_ok"] > 0 and out["import_pb_rejected"] > 0


@pytest.mark.timeout(300)
def test_product_codec_and_oracle_codec_agree_on_mutated_tuples(tmp_path):
    """tests/fuzz_codec_diff.cpp: the same verdict (count / SIB_EJSON / SIB_ESHAPE) and the same floats from
    csrc/sib_json.cpp and oracle/sib_oracle.c for mutated tuples, the same message text and Float.toString text for random
    floats. (Its first run found the oracle breaking exact decimal ties by rounding noise: tests/test_codec.py ties.)"""
    gxx, gcc = shutil.which("g++"), shutil.which("gcc")
    if not gxx or not gcc:
        pytest.skip("no gcc/g++")
    obj, exe = str(tmp_path / "oracle.o"), str(tmp_path / "fuzz_diff")
    r = subprocess.run([gcc, "-O1", "-g", "-c", os.path.join(ROOT, "oracle", "sib_oracle.c"), "-o", obj], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr[-2000:]
    cmd = [gxx, "-std=c++17", "-O1", "-g", "-I", os.path.join(ROOT, "include"), os.path.join(ROOT, "tests", "fuzz_codec_diff.cpp"),
           *[os.path.join(CSRC, f) for f in ("sib_json.cpp", "sib_pbimport.cpp", "sib_model.cpp")], obj, "-lpthread", "-lm", "-o", exe]
    r = subprocess.run(cmd, capture_output=True, text=True)
    assert r.returncode == 0, r.stderr[-2000:]
    r = subprocess.run([exe, "8"], capture_output=True, text=True)
    assert r.returncode == 0, (r.stdout[-500:], r.stderr[-3000:])
    out = json.loads(r.stdout.strip().splitlines()[-1])
    assert out["differences"] == 0 and out["both_accept"] > 1000 and out["both_reject"] > 1000
