"""ThreadSanitizer run of the library's HOST side (race detection): contexts, the sub-batch slot pipeline,
the combining dispatchers behind mm_chain_dp / mm_chain_dp_fpga / mm_chain_dp_bottom and mm2b200_chain_batch_multi.

csrc/chain_api.cu is compiled as plain C++ with g++ -fsanitize=thread against tests/tsan/cuda_runtime.h (an inert stand-in:
"device" memory is host memory, every async call completes at once) and tests/tsan/stub_kernels.cpp (trivial stand-ins for
the kernel launchers), then tests/tsan/stress.cpp hammers every entry point from ~30 threads.  No GPU involved; nothing of
this is part of the product."""
import os
import shutil
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
TS = os.path.join(ROOT, "tests", "tsan")


def test_host_side_is_race_free(tmp_path):
    if not shutil.which("g++"):
        pytest.skip("g++ not available")
    exe = str(tmp_path / "tsan_stress")
    cmd = ["g++", "-std=c++17", "-fsanitize=thread", "-O1", "-g", "-I", TS, "-I", os.path.join(ROOT, "include"),
           "-I", os.path.join(ROOT, "minimap2_chaindp_b200", "csrc"),
           "-x", "c++", os.path.join(ROOT, "minimap2_chaindp_b200", "csrc", "chain_api.cu"),
           "-x", "c++", os.path.join(TS, "stub_kernels.cpp"), os.path.join(TS, "stress.cpp"), "-o", exe, "-lpthread"]
    b = subprocess.run(cmd, capture_output=True, text=True)
    if b.returncode != 0 and "tsan" in (b.stderr or "").lower():
        pytest.skip("libtsan not available: " + b.stderr[-200:])
    assert b.returncode == 0, b.stderr[-2000:]
    env = dict(os.environ, TSAN_OPTIONS="halt_on_error=0 second_deadlock_stack=1", MM2B200_ON_ERROR="abort")
    r = subprocess.run([exe, "24", "400"], capture_output=True, text=True, env=env, timeout=600)
    out = r.stdout + r.stderr
    assert "ThreadSanitizer" not in out, out[-4000:]
    assert r.returncode == 0 and "stress: 0 mismatches" in out, out[-2000:]
