"""Coloured ``info/warn/error/success`` prints.

COINSTAC captures a computation's stdout as protocol JSON, so everything here goes to
stderr.  ``DINUNET_QUIET=1`` silences it (tests, benchmarks).
"""
import os
import sys

_C = {"info": "\033[96m", "warn": "\033[93m", "error": "\033[91m", "success": "\033[92m", "end": "\033[0m"}


def _emit(kind, msg, add_info=""):
    if os.environ.get("DINUNET_QUIET") == "1":
        return
    tty = sys.stderr.isatty()
    head = f"{_C[kind]}[{kind.upper()}]{_C['end']}" if tty else f"[{kind.upper()}]"
    print(f"{head} {add_info}{msg}", file=sys.stderr, flush=True)


def info(msg, verbose=True):
    if verbose:
        _emit("info", msg)


def warn(msg, verbose=True):
    if verbose:
        _emit("warn", msg)


def error(msg, verbose=True):
    if verbose:
        _emit("error", msg)


def success(msg, verbose=True):
    if verbose:
        _emit("success", msg)


def lazy_debug(x, add=0):
    """True on a sparse, geometrically thinning schedule — for progress prints."""
    import math
    x = int(x) + 1
    return x % int(math.log(x + 1) + add + 1) == 0
