"""mfcc_b200_librosa_run_static: the entry point that allocates nothing (its scratch is a file-scope
static __device__ array).  Called through the C ABI, two calls in a row, against the existing entry
point mfcc_b200_librosa_run with a caller-owned workspace.  Collected last on
purpose."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

torch = pytest.importorskip("torch")

ATOL, RTOL = 5e-3, 1e-4            # the tolerance of tests/test_librosa_mode.py


def test_librosa_run_static_matches_run_twice_in_a_row(wav):
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    import mfcc_b200
    from mfcc_b200 import _lib
    L = _lib.lib()
    dev = torch.device("cuda:0")
    op = mfcc_b200.LibrosaMFCC(sr=16000, hop_length=170, n_mfcc=32, device=dev)
    rng = np.random.default_rng(21)
    stream = C.c_void_p(torch.cuda.current_stream(dev).cuda_stream)

    def batch(n_clips, n, seed_off):
        rows = [wav[seed_off:seed_off + n]]
        for i in range(1, n_clips):
            rows.append(np.clip(rng.normal(0, 500.0 * (i + 1), n), -32768, 32767).astype(np.int16) if i % 3
                        else np.zeros(n, dtype=np.int16))
        return torch.from_numpy(np.stack(rows)).to(dev)

    # second call smaller than the first and with a silent clip first in line: stale maxima or rows
    # of call 1 left in the shared array would show
    calls = [batch(5, 170 * 200 + 7, 30000), batch(3, 170 * 57, 90000)]
    calls[1][0].zero_()
    with torch.cuda.device(dev):
        for x in calls:
            B, N = x.shape
            K = op.num_frames(N)
            want = op(x)                                        # mfcc_b200_librosa_run, torch-owned workspace
            got = torch.full((B, K, 32), float("nan"), dtype=torch.float32, device=dev)
            torch.cuda.synchronize()
            free0 = torch.cuda.mem_get_info(dev)[0]
            _lib.check(L.mfcc_b200_librosa_run_static(op._h, C.c_void_p(x.data_ptr()), x.stride(0), N, B, K,
                                                      C.c_void_p(got.data_ptr()), got.stride(0), stream))
            torch.cuda.synchronize()
            assert torch.cuda.mem_get_info(dev)[0] == free0     # the call took no device memory
            g, w = got.cpu().numpy().astype(np.float64), want.cpu().numpy().astype(np.float64)
            assert np.isfinite(g).all()
            assert (np.abs(g - w) <= ATOL + RTOL * np.abs(w)).all(), np.abs(g - w).max()
            assert np.array_equal(g, w)                         # same kernels: in fact bit for bit

        # beyond the static array: rejected with a message that names the limit, nothing launched
        K_big = _lib.LIBROSA_STATIC_BYTES // (128 * 4) + 1
        n0 = _lib.launch_count()
        with pytest.raises(_lib.MfccError, match="static scratch"):
            _lib.check(L.mfcc_b200_librosa_run_static(op._h, C.c_void_p(calls[0].data_ptr()), calls[0].stride(0),
                                                      170 * K_big, 1, K_big, C.c_void_p(calls[0].data_ptr()), K_big * 32,
                                                      stream))
        assert _lib.launch_count() == n0
        # the largest size that fits goes through the size check (5 clips of 6 540 frames would need real
        # input; here only the arithmetic of the limit: rows + rounded maxima == the array)
        assert 5 * 6540 * 128 * 4 + 256 <= _lib.LIBROSA_STATIC_BYTES < 5 * 6554 * 128 * 4 + 256
