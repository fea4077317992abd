"""Whole-program drop-in (INTEGRATION.md level 1): the reference's own minimap2 program, built from its unmodified
sources with chain.o left out and libmm2chain_b200.so linked in its place (oracle/Makefile `dropin`), must print the
golden PAF lines of tests/golden/paf.txt -- seeding, post-processing and output are the reference's, the chaining DP
(mm_chain_dp_fpga + mm_chain_dp_bottom, map.c:525,837) runs on the B200."""
import os
import subprocess

import pytest

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
BIN = os.path.join(ROOT, "oracle", "_ref", "minimap2_dropin")
DATA = os.path.join(ROOT, "oracle", "_ref", "test")


def _golden():
    out, name = {}, None
    for line in open(os.path.join(ROOT, "tests", "golden", "paf.txt")):
        if line.startswith("## "):
            name = line[3:].split(":")[0]
            out[name] = []
        elif line.strip():
            out[name].append("\t".join(line.rstrip("\n").split("\t")[:18]))
    return out


@pytest.mark.skipif(not os.path.exists(BIN), reason="oracle/_ref/minimap2_dropin not built (needs /root/reference at build time)")
def test_reference_program_with_gpu_chaining(tmp_path):
    both = tmp_path / "both.fa"
    both.write_bytes(open(os.path.join(DATA, "MT-human.fa"), "rb").read() + open(os.path.join(DATA, "MT-orang.fa"), "rb").read())
    runs = {
        "map-ont_MT": ["-x", "map-ont", f"{DATA}/MT-human.fa", f"{DATA}/MT-orang.fa"],
        "default_inv": [f"{DATA}/t-inv.fa", f"{DATA}/q-inv.fa"],
        "default_q2": [f"{DATA}/t2.fa", f"{DATA}/q2.fa"],
        "ava-ont_MT": ["-x", "ava-ont", str(both), str(both)],
    }
    want = _golden()
    for name, args in runs.items():
        p = subprocess.run([BIN, "-t", "12", *args], cwd=tmp_path, capture_output=True, text=True, timeout=300)
        assert p.returncode == 0, p.stderr[-2000:]
        got = ["\t".join(l.split("\t")[:18]) for l in p.stdout.splitlines() if l.strip()]
        assert sorted(got) == sorted(want[name]), f"{name}: PAF differs\n got {got}\nwant {want[name]}\n{p.stderr[-1500:]}"
