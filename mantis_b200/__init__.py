"""mantis_b200 -- B200 (sm_100a) implementation of mantis's per-frame localization hot path.

The product is ``libmantis_cuda.so`` (CUDA, C ABI, ``include/mantis_cuda.h``).  This package
is the thin Python host layer over that ABI used by the tests and ``bench.py``; the C++
hosts under ``mantis_b200/host`` are the drop-in mirrors of the reference's ``mantis3`` /
``mantis_server`` executables.  There is no CPU fallback: importing works anywhere, but
every compute call needs the built library and a CUDA device and raises otherwise.
"""
from .binding import (  # noqa: F401
    MEM_DEVICE,
    MEM_HOST,
    Config,
    Context,
    Frame,
    FrameLines,
    FrameQuads,
    FrameTrace,
    Result,
    RppResult,
    MantisError,
    PnpResult,
    build_library,
    library_path,
    load_library,
)
