#!/usr/bin/env python3
"""Generates tests/golden/tanh_y_metrics.npz: cell centres and the y metrics (deta_e, deta_d, deta_s) of a direction of
width 2 pi whose nodes follow the reference's `Tanh` map with the stretching of testcases/scalingTest/WeakScaling/base.json
(1.3843992415368547), periodic, for ny = 40 (parity tests), 512 and 1024 (weak-scaling bench: 512 cells per GPU, one or
two blocks in y).  Node placement: src/prometeo_grid.rg:123-230 (restated in htr_solver_b200/synthetic.py: grid_nodes);
metrics: the reference's own InitializeMetricTask arithmetic (src/prometeo_metric.cc:36-190, prometeo_metric.hpp:113-160)
evaluated through the oracle (oracle/_ref).  Needs /root/reference; the fixture is committed, bench.py only reads it.
    python tests/golden/gen_tanh_metrics.py"""
import os, sys
import numpy as np
HERE = os.path.dirname(os.path.abspath(__file__))
sys.path[:0] = [os.path.dirname(os.path.dirname(HERE)), os.path.dirname(HERE)]
from oracle_lib import Oracle, tanh_y_metrics
from htr_solver_b200 import synthetic

STRETCHING = 1.3843992415368547
WIDTH = 2 * np.pi


def main():
    o = Oracle("ConstPropMix")
    out = {"stretching": STRETCHING, "width": WIDTH}
    for ny in (40, 512, 1024):
        m = tanh_y_metrics(o, ny, STRETCHING, WIDTH)
        for k, v in m.items():
            out[f"ny{ny}_{k}"] = v
        print(ny, "cell size ratio max/min", (1 / m["deta_e"]).max() / (1 / m["deta_e"]).min())
    np.savez(os.path.join(HERE, "tanh_y_metrics.npz"), **out)


if __name__ == "__main__":
    main()
