#!/usr/bin/env python
"""The reference's own acceptance rule for its CUDA build, as a command-line tool.

colab-runner.ipynb (cell 4, `compare_outputs_with_tolerance`) compares the stdout of `serial` and
`parallel` token by token: numeric tokens must agree within tol = 1e-4 -- absolutely when the
reference value is <= 1 in magnitude, relatively otherwise -- every other token must be equal.

  python tools/compare_outputs.py out_a.txt out_b.txt [tol]      -> exit 0 if equivalent
"""
import sys


def tokens(path):
    return open(path).read().split()


def equivalent(a, b, tol=1e-4):
    ta, tb = tokens(a), tokens(b)
    if len(ta) != len(tb):
        return False, f"token count {len(ta)} vs {len(tb)}"
    for k, (x, y) in enumerate(zip(ta, tb)):
        try:
            fx, fy = float(x), float(y)
        except ValueError:
            if x != y:
                return False, f"token {k}: {x!r} vs {y!r}"
            continue
        err = abs(fx - fy) if abs(fx) <= 1.0 else abs(fx - fy) / abs(fx)
        if err > tol:
            return False, f"token {k}: {fx} vs {fy} (error {err:.3g} > {tol})"
    return True, f"{len(ta)} tokens agree within {tol}"


if __name__ == "__main__":
    tol = float(sys.argv[3]) if len(sys.argv) > 3 else 1e-4
    ok, msg = equivalent(sys.argv[1], sys.argv[2], tol)
    print(("EQUIVALENT: " if ok else "DIFFERENT: ") + msg)
    sys.exit(0 if ok else 1)
