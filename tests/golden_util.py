"""Loader for the golden vectors under tests/golden/ (made by tests/golden/make_golden.sh from the reference itself)."""
import glob
import os
import re

import numpy as np

GOLDEN_DIR = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
PARAM_NAMES = ("max_dist_x", "max_dist_y", "bw", "max_skip", "min_cnt", "min_sc", "is_cdna", "n_segs")


def golden_calls():
    """-> list of dicts: name, params (dict), a, rec, u, b -- one per mm_chain_dp call the reference made."""
    out = []
    for path in sorted(glob.glob(os.path.join(GOLDEN_DIR, "*.npz"))):
        if os.path.basename(path).startswith("seed_"):     # fixtures of the seeding stage (tests/test_seed_oracle.py)
            continue
        z = np.load(path)
        for k in range(int(z["n_calls"][0])):
            out.append(dict(name=f"{os.path.basename(path)[:-4]}#{k}",
                            params=dict(zip(PARAM_NAMES, (int(x) for x in z[f"c{k}_params"]))),
                            a=z[f"c{k}_a"], rec=z[f"c{k}_rec"], u=z[f"c{k}_u"], b=z[f"c{k}_b"]))
    return out


def paf_pins():
    """(command name, chain score s1:i, anchor count cm:i) for every PAF line the reference printed."""
    pins, name = [], None
    for line in open(os.path.join(GOLDEN_DIR, "paf.txt")):
        if line.startswith("## "):
            name = line[3:].split(":")[0]
            continue
        s1 = re.search(r"s1:i:(\d+)", line)
        cm = re.search(r"cm:i:(\d+)", line)
        if s1 and cm:
            pins.append((name, int(s1.group(1)), int(cm.group(1))))
    return pins
