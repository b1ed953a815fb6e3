"""GPU (-m gpu): log-mel kernel against the numpy oracle (same seeded inputs), the committed golden vectors,
edge cases (silence, pure tone, ragged / very short clips, 401-frame clip, edge padding) and, at
BASELINE.json's full batch size, size-independent properties (batch independence, gain linearity)."""
import os

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

from oracle import logmel_np as OL  # noqa: E402

# tolerances: 1e-4 relative on log values for bins within 80 dB of the clip's peak, and 1e-5 of the
# peak in the linear domain on every bin (log(x + 1e-6) is ill-conditioned below that; see oracle.logmel_parity)
TOL_LOG, TOL_LIN = 1e-4, 1e-5


def _cfg():
    from asf_b200.config.defaults import epic_slowfast_cfg
    return epic_slowfast_cfg()


def _gpu_logmel(w, num_frames=None, impl=None):
    from asf_b200.datasets.audio_loader_epic import logmel_cuda
    x = torch.from_numpy(np.ascontiguousarray(w)).cuda()
    if x.dim() == 1:
        x = x[None]
    return logmel_cuda(x, num_frames=num_frames, impl=impl).cpu().numpy()


CASES = ["noise0", "noise1", "sine1k", "zeros", "len48000", "len23975", "len5000", "len119"]


def _inputs():
    wav = OL.synthetic_waveforms(2, 47975, 1234)
    t = np.arange(47975) / 24000.0
    return {
        "noise0": wav[0], "noise1": wav[1],
        "sine1k": (0.5 * np.sin(2 * np.pi * 1000.0 * t)).astype(np.float32),
        "zeros": np.zeros(47975, np.float32),
        "len48000": OL.synthetic_waveforms(1, 48000, 99)[0],
        "len23975": OL.synthetic_waveforms(1, 23975, 98)[0],
        "len5000": OL.synthetic_waveforms(1, 5000, 97)[0],
        "len119": OL.synthetic_waveforms(1, 119, 96)[0],
    }


@pytest.mark.parametrize("impl", ["tc", "simt"])  # tcgen05 3xTF32 kernel (the default) / fp32 CUDA-core kernel
@pytest.mark.parametrize("name", CASES)
def test_logmel_matches_oracle_and_golden(golden_dir, name, impl):
    w = _inputs()[name]
    out = _gpu_logmel(w, impl=impl)[0, 0]
    ref = OL.log_specgram(w)
    gold = np.load(os.path.join(golden_dir, "logmel.npz"))[name]
    assert out.shape == ref.shape == gold.shape
    for r in (ref, gold):
        rel_log, rel_lin, _ = OL.logmel_parity(r, out)
        assert rel_log <= TOL_LOG and rel_lin <= TOL_LIN, (name, rel_log, rel_lin)


def test_reference_signature_host_buffers():
    from asf_b200.datasets.audio_loader_epic import _log_specgram
    cfg = _cfg()
    w = OL.synthetic_waveforms(1, 47975, 5)[0]
    out = _log_specgram(cfg, w, window_size=cfg.AUDIO_DATA.WINDOW_LENGTH, step_size=cfg.AUDIO_DATA.HOP_LENGTH)
    assert isinstance(out, np.ndarray) and out.shape == (400, 128) and out.dtype == np.float32
    rel_log, rel_lin, _ = OL.logmel_parity(OL.log_specgram(w), out)
    assert rel_log <= TOL_LOG and rel_lin <= TOL_LIN
    # other window / hop ratio exercises the EPIC hop quirk (hop = win - hop)
    out2 = _log_specgram(cfg, w[:20000], window_size=10.0, step_size=2.5)
    ref2 = OL.log_specgram(w[:20000], window_ms=10.0, hop_ms=2.5)
    assert out2.shape == ref2.shape
    assert OL.logmel_parity(ref2, out2)[0] <= TOL_LOG


def test_extract_sound_feature_and_batched_pack():
    from types import SimpleNamespace
    from asf_b200.datasets.audio_loader_epic import _extract_sound_feature, logmel_batch
    cfg = _cfg()
    w = OL.synthetic_waveforms(1, 60000, 6)[0]
    rec = SimpleNamespace(num_audio_samples=60000, start_audio_sample=0, end_audio_sample=60000)
    spec = _extract_sound_feature(cfg, w, rec, 100, 100 + 47975)
    assert spec.shape == (1, 400, 128)
    ref = OL.extract_sound_feature(w[100:100 + 47975])
    assert OL.logmel_parity(ref[0], spec[0].numpy())[0] <= TOL_LOG
    short = SimpleNamespace(num_audio_samples=5000, start_audio_sample=10, end_audio_sample=5010)
    sp = _extract_sound_feature(cfg, w, short, 0, 47975)
    assert sp.shape == (1, 400, 128) and torch.equal(sp[0, 41], sp[0, 399])  # edge-padded
    assert OL.logmel_parity(OL.extract_sound_feature(w[10:5010])[0], sp[0].numpy())[0] <= TOL_LOG
    # batched: [slow, fast] with the reference's linspace gather (bit-exact copy of fast rows)
    wav = torch.from_numpy(OL.synthetic_waveforms(3, 47975, 7)).cuda()
    slow, fast = logmel_batch(cfg, wav)
    assert slow.shape == (3, 1, 100, 128) and fast.shape == (3, 1, 400, 128)
    idx = torch.from_numpy(OL.slow_indices(400, 4)).cuda()
    assert torch.equal(slow, fast[:, :, idx, :])


def test_tc_tiles_straddle_clips_and_edge_padding():
    """The tensor-core kernel tiles the flattened (clip, frame) space in 128-row tiles: clips of 42 frames put three
    clip boundaries inside one tile; num_frames > T adds the np.pad(..., "edge") rows; both must agree with the
    per-clip oracle and with the CUDA-core kernel."""
    wav = OL.synthetic_waveforms(37, 5000, 21)
    tc = _gpu_logmel(wav, num_frames=50, impl="tc")
    simt = _gpu_logmel(wav, num_frames=50, impl="simt")
    assert tc.shape == simt.shape == (37, 1, 50, 128)
    for b in (0, 1, 2, 3, 17, 36):
        ref = OL.extract_sound_feature(wav[b], num_frames=50)[0]
        for out in (tc, simt):
            rel_log, rel_lin, _ = OL.logmel_parity(ref, out[b, 0])
            assert rel_log <= TOL_LOG and rel_lin <= TOL_LIN, (b, rel_log, rel_lin)
    assert np.array_equal(tc[:, 0, 41], tc[:, 0, 49])
    # loud / quiet / DC-offset clips in one batch: per-clip independence of the shared accumulators
    mix = OL.synthetic_waveforms(5, 47975, 22)
    mix[1] *= 1e-3
    mix[2] = 0.0
    mix[3] += 0.25
    out = _gpu_logmel(mix, impl="tc")
    for b in range(5):
        rel_log, rel_lin, _ = OL.logmel_parity(OL.log_specgram(mix[b]), out[b, 0])
        assert rel_log <= TOL_LOG and rel_lin <= TOL_LIN, (b, rel_log, rel_lin)


def test_full_batch_properties():
    """BASELINE config 2 batch (256 clips): batch independence and gain linearity of the linear-domain mel."""
    from asf_b200.datasets.audio_loader_epic import logmel_cuda
    wav = torch.from_numpy(OL.synthetic_waveforms(256, 47975, 11)).cuda()
    full = logmel_cuda(wav)
    assert full.shape == (256, 1, 400, 128) and torch.isfinite(full).all()
    single = logmel_cuda(wav[17:18])
    assert torch.equal(full[17:18], single)
    half = logmel_cuda(0.5 * wav[:8])
    lin_full = torch.exp(full[:8]) - 1e-6
    lin_half = torch.exp(half) - 1e-6
    assert float(((lin_half - 0.5 * lin_full).abs() / lin_full.abs().max()).max()) < 1e-5
    ref = OL.log_specgram(wav[255].cpu().numpy())
    assert OL.logmel_parity(ref, full[255, 0].cpu().numpy())[0] <= TOL_LOG
