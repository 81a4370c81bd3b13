"""avg.sh (avg.sh:1-15): for every *.txt in the current directory (or the files given), grep a
pattern (default `gather`), split the matching lines on ':' and print the mean of field 2 --
i.e. the rank-average of a `TIME <phase> :` line."""
from __future__ import annotations

import glob
import sys


def average(path: str, pattern: str) -> tuple:
    vals = []
    with open(path, errors="replace") as f:
        for line in f:
            if pattern in line and ":" in line:
                try:
                    vals.append(float(line.split(":")[1].split()[0]))
                except (ValueError, IndexError):
                    continue
    return (sum(vals) / len(vals) if vals else float("nan")), len(vals)


def main(argv=None) -> int:
    argv = list(sys.argv[1:] if argv is None else argv)
    pattern = argv[0] if argv else "gather"
    files = argv[1:] or sorted(glob.glob("*.txt"))
    print("PATTERN=%s" % pattern)
    for path in files:
        mean, n = average(path, pattern)
        print("%s %s" % (path, "%.6g" % mean if n else "-nan"))
    return 0


if __name__ == "__main__":
    sys.exit(main())
