"""daxpy_nvtx (daxpy_nvtx.cu:38-108): daxpy with NVTX ranges copyInput / cublasDaxpy / copyOutput
inside a cudaProfilerStart/Stop window."""
import sys

from .daxpy import run


def main(argv=None) -> int:
    return run(argv, with_nvtx=True)


if __name__ == "__main__":
    sys.exit(main())
