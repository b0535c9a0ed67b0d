#!/usr/bin/env python3
"""The reference's scripts/compare_hdf.py for field dumps of this repository.

    python tools/compare_fields.py A B [--solver-tol]

A and B are two dumps of the same fields: .npz files (numpy.savez of what Fluid.download() returns, or of an oracle field
dictionary) or, where h5py is importable, HDF5 restart tiles as the reference writes them.  For every field it prints the
reference's line (scripts/compare_hdf.py:17-26)

    <field>: max diff = <max |A - B|> at <index>

and asserts that both files hold the same fields with the same shapes.  With --solver-tol it also applies the rule of the
reference's solver tests (solverTests/testAll.py:23-36: |A - B| <= max(1e-10 |B|, 1e-12) point-wise, B = the reference file)
and exits 1 on the first field that breaks it, with the reference's error line.
"""
import argparse
import sys

import numpy as np


def load(path):
    if path.endswith((".hdf", ".h5", ".hdf5")):
        import h5py      # absent on the build image; present where real HTR restarts are
        with h5py.File(path, "r") as f:
            return {k: f[k][:] for k in f}
    with np.load(path) as z:
        return {k: z[k] for k in z.files}


def compare(a, b, out=sys.stdout):
    """The lines of compare_hdf.py; returns {field: (max diff, location)}."""
    for fld in a:
        assert fld in b and a[fld].shape == b[fld].shape, fld
    for fld in b:
        assert fld in a and a[fld].shape == b[fld].shape, fld
    res = {}
    for fld in a:
        diff = np.absolute(a[fld].astype(np.float64) - b[fld].astype(np.float64))
        max_diff = np.max(diff) if diff.size else 0.0
        max_loc = tuple(int(i) for i in np.unravel_index(np.argmax(diff), diff.shape)) if diff.size else ()
        print('%s: max diff = %s at %s' % (fld, max_diff, max_loc), file=out)      # plain ints: numpy 2 would print np.int64(..)
        res[fld] = (float(max_diff), max_loc)
    return res


def check_solution(sol, ref, tol=1e-10, out=sys.stdout):
    """checkSolutionFile of solverTests/testAll.py: True if every field of `ref` is matched within max(tol |ref|, 1e-12)."""
    for fld in ref:
        f = np.asarray(sol[fld], dtype=np.float64)
        r = np.asarray(ref[fld], dtype=np.float64)
        diff = np.absolute(f - r)
        if (diff > np.maximum(tol*abs(r), 1e-12)).any():
            max_loc = tuple(int(i) for i in np.unravel_index(np.argmax(diff), diff.shape))
            ref_val = r[max_loc]
            print("Error on field: {}, delta = {}, location = {}, tol = {}, expected = {}".format(
                  fld, np.max(diff), max_loc, max(tol*abs(ref_val), 1e-12), ref_val), file=out)
            return False
    return True


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("f1")
    ap.add_argument("f2", help="the reference file for --solver-tol")
    ap.add_argument("--solver-tol", action="store_true")
    args = ap.parse_args()
    a, b = load(args.f1), load(args.f2)
    compare(a, b)
    if args.solver_tol and not check_solution(a, b):
        sys.exit(1)


if __name__ == "__main__":
    main()
