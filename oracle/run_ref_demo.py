#!/usr/bin/env python
"""Test infrastructure: execute one of the reference's UNMODIFIED demo scripts from the byte-compiled archive
oracle/_ref/pyref.zip (oracle/build_ref_solver.py) as if it had been started as ``python demo_x.py ...`` in the
current directory:

    cd <scratch>/demo && PYTHONPATH=<repo>:<repo>/compat python <repo>/oracle/run_ref_demo.py demo_2d_LBM --total_steps 20

``__name__`` is "__main__", ``__file__`` is <cwd>/<name>.py (the demos write their frames next to ``__file__``),
``sys.argv`` and ``sys.path[0]`` are what the interpreter would have set.  Which ``src.LBM`` the script gets is decided
by PYTHONPATH alone: tests/test_gpu_demos.py puts this repo's drop-in shim there."""
import os
import sys

HERE = os.path.dirname(os.path.abspath(__file__))


def main():
    name = sys.argv[1]
    sys.path.insert(0, HERE)
    try:
        import build_ref_solver
        code = build_ref_solver.demo_code(name)
    finally:
        sys.path.remove(HERE)
    script = os.path.join(os.getcwd(), name + ".py")
    sys.argv = [script] + sys.argv[2:]
    sys.path[0:0] = [os.getcwd()]
    exec(code, {"__name__": "__main__", "__file__": script, "__builtins__": __builtins__})


if __name__ == "__main__":
    main()
