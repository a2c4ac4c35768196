"""honk_frontend_mfcc_noalloc (include/honk_b200.h): the allocation-free variant of the int16 front end, whose scratch is
a file-scope static device array.  No counterpart in the reference.

Checked through the C ABI (ctypes) against the existing entry point honk_frontend_mfcc, on two calls in a row (the array
keeps the previous call's contents), within the front-end tolerance the other GPU tests use (2e-4; the two entry points run
the same kernels, so they are in fact bit-identical).  Collected last on purpose (file name).
"""
import numpy as np
import pytest
import torch

from honk_b200 import synth

pytestmark = pytest.mark.gpu
ATOL = 2e-4
MAX_FRAMES = 6464                      # HONK_NOALLOC_MAX_FRAMES


def _noalloc(pre, pcm, out):
    from honk_b200 import _native as N
    rc = N.lib().honk_frontend_mfcc_noalloc(pre.handle, pcm.data_ptr(), pcm.shape[0], pcm.shape[1], pcm.stride(0), out.data_ptr(),
                                            N.stream_ptr(pcm.device))
    N.check(rc)
    return out


def test_noalloc_matches_existing_entry_point_twice():
    import honk_b200 as hb
    pre = hb.AudioPreprocessor()
    assert pre.native_path == "tcgen05-i8-limbs"
    first = torch.from_numpy(synth.make_clips(64, start=31000)).cuda()             # the maximum: 64 x 101 = 6464 frames
    second = torch.from_numpy(synth.make_clips(5, start=32000)).cuda()             # fewer frames: stale rows stay behind it
    # outputs are allocated by the caller BEFORE the calls; the entry point itself allocates nothing
    out1 = torch.full((64, 101, 40), float("nan"), device="cuda")
    out2 = torch.full((5, 101, 40), float("nan"), device="cuda")
    torch.cuda.synchronize()
    _noalloc(pre, first, out1)
    _noalloc(pre, second, out2)
    torch.cuda.synchronize()
    ref1, ref2 = pre.compute_mfccs_batch(first), pre.compute_mfccs_batch(second)   # honk_frontend_mfcc with a caller workspace
    for got, ref in ((out1, ref1), (out2, ref2)):
        assert not torch.isnan(got).any()
        assert (got - ref).abs().max().item() <= ATOL
    # and once more in the other order, on the same (now dirty) scratch
    out2.fill_(float("nan"))
    out1.fill_(float("nan"))
    _noalloc(pre, second, out2)
    _noalloc(pre, first, out1)
    torch.cuda.synchronize()
    assert (out1 - ref1).abs().max().item() <= ATOL and (out2 - ref2).abs().max().item() <= ATOL
    # strided rows (sliding windows over one buffer) work like in the existing entry point
    buf = torch.from_numpy(np.concatenate(list(synth.make_clips(2, start=33000)))).cuda()
    win = buf.as_strided((3, 16000), (8000, 1))
    out3 = torch.empty((3, 101, 40), device="cuda")
    _noalloc(pre, win, out3)
    assert (out3 - pre.compute_mfccs_batch(win)).abs().max().item() <= ATOL


def test_noalloc_rejects_sizes_beyond_the_static_scratch():
    import honk_b200 as hb
    from honk_b200 import _native as N
    pre = hb.AudioPreprocessor()
    big = torch.zeros((65, 16000), dtype=torch.int16, device="cuda")               # 65 x 101 frames > 6464
    out = torch.empty((65, 101, 40), device="cuda")
    with pytest.raises(ValueError, match="exceed the static scratch"):
        _noalloc(pre, big, out)
    long = torch.zeros((1, 160 * MAX_FRAMES), dtype=torch.int16, device="cuda")     # one clip of 6465 frames
    with pytest.raises(ValueError, match="exceed the static scratch"):
        _noalloc(pre, long, torch.empty((1, MAX_FRAMES + 1, 40), device="cuda"))
    other = hb.AudioPreprocessor(n_mels=32, n_dct_filters=20)                       # not the default geometry
    with pytest.raises(NotImplementedError):
        _noalloc(other, big[:1], torch.empty((1, 101, 20), device="cuda"))
    assert N.lib().honk_frontend_mfcc_noalloc(pre.handle, None, 0, 16000, 16000, None, None) == 0   # empty batch: nothing to do
    _noalloc(pre, big[:2], out[:2])                                                  # the handle still works after the refusals
    torch.cuda.synchronize()
    assert torch.all(out[:2] == 0)                                                   # silence stays exactly zero (manage_audio.py:39)
