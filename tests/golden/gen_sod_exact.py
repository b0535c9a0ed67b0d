#!/usr/bin/env python3
"""Copies the reference's exact solution of the Sod shock tube at t = 0.2 (testcases/SodProblem/exact.txt: 1001 rows of
x in [-1, 1], rho, u, p -- data, read by testcases/SodProblem/postProc.py:17-39) to tests/golden/sod_exact.txt so that
the GPU test of BASELINE config C1 can read it where /root/reference does not exist.
    python tests/golden/gen_sod_exact.py"""
import os, shutil
HERE = os.path.dirname(os.path.abspath(__file__))
shutil.copyfile("/root/reference/testcases/SodProblem/exact.txt", os.path.join(HERE, "sod_exact.txt"))
print("wrote", os.path.join(HERE, "sod_exact.txt"))
