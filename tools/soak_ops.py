#!/usr/bin/env python
"""Soak run (GPU box): the seeded random-shape checks of tests/test_gpu_fuzz_shapes.py with many more seeds than the
suite carries (K1 with every channel / view / lane-width class incl. gradients, K2, K4, K3), each against the CPU oracle.
Prints one line per failure and a summary; exit code = failures (capped).

    python tools/soak_ops.py [seeds_per_op=150] [seed0=100]"""
import os
import sys
import traceback

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import test_gpu_fuzz_shapes as fz  # noqa: E402


def main():
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 150
    s0 = int(sys.argv[2]) if len(sys.argv) > 2 else 100
    fails = 0
    for name in ("test_unproject_random_shape", "test_soft_argmax_3d_random_shape", "test_soft_argmax_2d_random_shape",
                 "test_dlt_random_shape"):
        fn, bad = getattr(fz, name), 0
        for seed in range(s0, s0 + n):
            try:
                fn(seed)
            except AssertionError:
                bad += 1
                print("FAIL %s seed %d: %s" % (name, seed, traceback.format_exc().strip().splitlines()[-1][:200]))
        print("%s: %d seeds, %d failures" % (name, n, bad))
        fails += bad
    return fails


if __name__ == "__main__":
    sys.exit(min(main(), 100))
