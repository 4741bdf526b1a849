"""stitch_device_digest_noalloc (GPU): the digest entry point that allocates nothing -- scratch in a file-scope
__device__ array (imagestitching_b200/csrc/noalloc.cu) -- against stitch_device_digest, through the C ABI.

No counterpart in the reference.  Integer arithmetic, so the bar is
bit-exact.  Two calls in a row on different buffers (the scratch keeps the first call's contents), then the size limit.
Collected last on purpose (`zz`)."""
import ctypes as C

import numpy as np
import pytest

from tests import util as U


@pytest.mark.gpu
def test_noalloc_digest_equals_digest_twice_in_a_row():
    from imagestitching_b200 import StitchContext
    from imagestitching_b200 import _lib as L
    ctx = StitchContext(0)
    lib = ctx.lib
    stream = lib.stitch_ctx_stream(ctx.handle)
    rng = np.random.default_rng(7)
    # set-up (may allocate): input buffers and the 8-byte device word for the result
    out = C.c_void_p()
    assert lib.stitch_device_alloc(ctx.handle, 8, C.byref(out)) == 0
    err = C.create_string_buffer(256)
    # a 3.1 MB buffer (13 partials, odd tail bytes), then a SMALLER one (2 partials: 11 stale slots stay in the scratch),
    # then a misaligned view and a tiny one
    for nbytes, offset in ((3_100_003, 0), (300_001, 0), (1_000_000, 1), (3, 0)):
        host = rng.integers(0, 256, nbytes + offset, dtype=np.uint8)
        buf = C.c_void_p()
        assert lib.stitch_device_alloc(ctx.handle, nbytes + offset, C.byref(buf)) == 0
        assert lib.stitch_memcpy_h2d(ctx.handle, buf, host.ctypes.data, nbytes + offset) == 0
        ptr = buf.value + offset
        want = C.c_uint64(0)
        assert lib.stitch_device_digest(ctx.handle, ptr, nbytes, C.byref(want)) == 0
        for _ in range(2):   # two calls in a row on the same input give the same answer
            st = lib.stitch_device_digest_noalloc(0, stream, ptr, nbytes, out, err, len(err))
            assert st == 0, err.value
            got = np.zeros(1, np.uint64)
            assert lib.stitch_memcpy_d2h(ctx.handle, got.ctypes.data, out, 8) == 0   # same stream: ordered after the kernels
            assert int(got[0]) == int(want.value) == U.digest(host[offset:]), (nbytes, offset)
        lib.stitch_device_free(ctx.handle, buf)
    # beyond the static scratch: refused with a message, nothing launched
    st = lib.stitch_device_digest_noalloc(0, stream, out, L.NOALLOC_DIGEST_MAX_BYTES + 1, out, err, len(err))
    assert st == L.E_BAD_ARG and b"STITCH_NOALLOC_DIGEST_MAX_BYTES" in err.value and str(L.NOALLOC_DIGEST_MAX_BYTES + 1).encode() in err.value
    assert lib.stitch_device_digest_noalloc(0, stream, None, 16, out, None, 0) == L.E_BAD_ARG
    lib.stitch_device_free(ctx.handle, out)
    ctx.close()


def test_noalloc_source_never_allocates():
    """The entry point's translation unit names no allocating call."""
    import os
    import re
    src = open(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "imagestitching_b200", "csrc", "noalloc.cu"),
               encoding="utf-8").read()
    code = re.sub(r"//[^\n]*", "", src)
    for banned in ("cudaMalloc", "cudaFree", "cudaHostAlloc", "cudaMallocHost", "cudaMallocAsync", "new ", "malloc(", "std::vector"):
        assert banned not in code, banned
    assert "cudaGetSymbolAddress" in code and "static __device__ __align__(256)" in code
