"""Shared plumbing of the headless demo ports: every ``demo_*.py`` here mirrors the script of the same name under the
reference's ``examples/`` (Python 2 + matplotlib + interactive ``raw_input``), keeps its model, sizes, seed and loop,
and replaces the plots by the traces the plots were drawn from (returned, printed as a short summary, optionally
written to an .npz).  No graphics stack is needed or used ("unavailable on this box")."""
import argparse
import os
import sys

import numpy as np

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))


def seed_everything(seed):
    if seed is None:
        seed = np.random.randint(2 ** 32)
    print("Setting seed to ", seed)
    np.random.seed(seed)
    return seed


def report(name, traces, out=None):
    """Print first / last / best of every trace; write them to ``out`` (.npz) when given."""
    for key, val in traces.items():
        v = np.asarray(val, dtype=float)
        if v.ndim == 1 and v.size:
            print("%s: %s  first %.4f  last %.4f  max %.4f  (%d values)" % (name, key, v[0], v[-1], np.nanmax(v), v.size))
    if out:
        np.savez(out, **{k: np.asarray(v) for k, v in traces.items()})
        print("wrote", out)
    return traces


def cli(demo, default_iters, default_seed=None, iters_name="N_samples"):
    ap = argparse.ArgumentParser(description=demo.__doc__)
    ap.add_argument("--seed", type=int, default=default_seed)
    ap.add_argument("--iters", type=int, default=default_iters, help="iterations (reference default: %d)" % default_iters)
    ap.add_argument("--out", default=None, help="write the traces to this .npz")
    ap.add_argument("--quiet", action="store_true")
    a = ap.parse_args()
    traces = demo(seed=a.seed, verbose=not a.quiet, **{iters_name: a.iters})
    report(demo.__module__ if demo.__module__ != "__main__" else os.path.basename(sys.argv[0]), traces, a.out)
