"""mpienv (mpienv.f90:1-35): every rank prints `rank <r> MEMORY_PER_CORE=<int>` read from the
environment -- a debug aid for launchers that swallow environment variables."""
from __future__ import annotations

import os
import sys

from ..utils import report
from . import _common


def main(argv=None) -> int:
    ap = _common.base_parser(__doc__)
    args = ap.parse_args(argv)
    rt, out, impl = _common.start(args)
    raw = os.environ.get("MEMORY_PER_CORE", "")
    try:
        value = int(raw)
    except ValueError:
        value = 0  # the Fortran program reads an unset variable as 0
    out(report.mpienv_line(rt.rank, value))
    _common.finish(rt)
    return 0


if __name__ == "__main__":
    sys.exit(main())
