"""afe_fbank_batch_static: the no-allocation variant of the fbank entry point (static device scratch).

Collected last on purpose (file name).  Checked through the C ABI against
afe_fbank_batch on the same inputs -- bit-identical (same kernels, same in-order reductions), on two calls in a
row (the static array keeps the first call's contents), in a fresh process or after the other tests."""
import ctypes as C

import numpy as np
import pytest

from audio_processor_b200 import _native


def _case(torch, dev, seed, lens):
    g = torch.Generator(device="cpu"); g.manual_seed(seed)
    waves = [((torch.rand(n, generator=g) - 0.5) * 0.8).to(dev) for n in lens]
    return waves


def _run(fe, torch, waves, lens, group, static):
    lib = fe.lib
    B = len(lens)
    lens_c = _native.i64_array(lens)
    ng = (B + group - 1) // group
    tps = (C.c_int64 * ng)(); rows = C.c_int64()
    lib.check(lib.afe_fbank_layout(lens_c, B, group, tps, C.byref(rows)))
    feat = torch.full((rows.value // 2, 160), float("nan"), dtype=torch.float32, device=fe.device)
    mask = torch.full((rows.value // 2,), -7, dtype=torch.int32, device=fe.device)
    ptrs = _native.ptr_array([w.data_ptr() for w in waves])
    stream = C.c_void_p(torch.cuda.current_stream(fe.device).cuda_stream)
    if static:
        code = lib.afe_fbank_batch_static(fe.ctx, stream, ptrs, lens_c, B, group, 0.0, _native.PRECISION_F64, 0,
                                          C.c_void_p(feat.data_ptr()), C.c_void_p(mask.data_ptr()))
    else:
        ws = torch.empty(int(lib.afe_fbank_workspace_bytes(lens_c, B)), dtype=torch.uint8, device=fe.device)
        code = lib.afe_fbank_batch(fe.ctx, stream, ptrs, lens_c, B, group, 0.0, _native.PRECISION_F64, 0,
                                   C.c_void_p(feat.data_ptr()), C.c_void_p(mask.data_ptr()),
                                   C.c_void_p(ws.data_ptr()), ws.numel())
    lib.check(code)
    torch.cuda.synchronize(fe.device)
    return feat.cpu().numpy(), mask.cpu().numpy()


@pytest.mark.gpu
def test_fbank_static_scratch_matches_workspace_entry_point_twice():
    import torch
    from audio_processor_b200 import runtime
    fe = runtime.get_frontend(0)
    # call 1: ragged utterances incl. one too short for a frame and one spanning several blocks' tile ranges;
    # call 2: a different, smaller batch, so stale scratch of call 1 would show
    for seed, lens, group in ((11, [400_000, 399, 16_000, 123_457, 400, 720, 48_000, 250_001], 4),
                              (12, [30_000, 560, 80_000], 16)):
        waves = _case(torch, fe.device, seed, lens)
        f_ref, m_ref = _run(fe, torch, waves, lens, group, static=False)
        f_new, m_new = _run(fe, torch, waves, lens, group, static=True)
        assert np.array_equal(m_ref, m_new)
        assert np.array_equal(f_ref, f_new, equal_nan=True)
        assert np.isfinite(f_new[m_new == 1]).all()


@pytest.mark.gpu
def test_fbank_static_scratch_rejects_oversize_calls():
    import torch
    from audio_processor_b200 import runtime
    fe = runtime.get_frontend(0)
    lib = fe.lib
    stream = C.c_void_p(torch.cuda.current_stream(fe.device).cuda_stream)
    one = torch.zeros(1, dtype=torch.float32, device=fe.device)
    # more waveforms than AFE_FBANK_STATIC_MAX_BATCH
    n = _native.FBANK_STATIC_MAX_BATCH + 1
    code = lib.afe_fbank_batch_static(fe.ctx, stream, _native.ptr_array([one.data_ptr()] * n), _native.i64_array([1] * n),
                                      n, 0, 0.0, _native.PRECISION_F64, 0, None, None)
    assert code == _native.AFE_ERR_WORKSPACE and b"maximum" in lib.afe_last_error()
    # more scratch than the static array holds (nothing is read before the check)
    lens = [16000 * 3600]
    assert int(lib.afe_fbank_workspace_bytes(_native.i64_array(lens), 1)) > _native.FBANK_STATIC_SCRATCH_BYTES
    code = lib.afe_fbank_batch_static(fe.ctx, stream, _native.ptr_array([one.data_ptr()]), _native.i64_array(lens),
                                      1, 0, 0.0, _native.PRECISION_F64, 0, None, None)
    assert code == _native.AFE_ERR_WORKSPACE and b"static array" in lib.afe_last_error()
