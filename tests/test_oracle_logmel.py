"""CPU: the numpy log-mel oracle against its committed golden vectors, against an independent
implementation (torchaudio) and basic invariants of the restated librosa algorithm."""
import os

import numpy as np
import pytest
import torch

from oracle import logmel_np as L


@pytest.fixture(scope="module")
def gold(golden_dir):
    return np.load(os.path.join(golden_dir, "logmel.npz"))


def _cases():
    wav = L.synthetic_waveforms(2, 47975, 1234)
    t = np.arange(47975) / 24000.0
    return {
        "noise0": wav[0], "noise1": wav[1],
        "sine1k": (0.5 * np.sin(2 * np.pi * 1000.0 * t)).astype(np.float32),
        "zeros": np.zeros(47975, np.float32),
        "len48000": L.synthetic_waveforms(1, 48000, 99)[0],
        "len23975": L.synthetic_waveforms(1, 23975, 98)[0],
        "len5000": L.synthetic_waveforms(1, 5000, 97)[0],
        "len119": L.synthetic_waveforms(1, 119, 96)[0],
    }


@pytest.mark.parametrize("name", ["noise0", "noise1", "sine1k", "zeros", "len48000", "len23975", "len5000", "len119"])
def test_oracle_matches_golden(gold, name):
    out = L.log_specgram(_cases()[name])
    ref = gold[name]
    assert out.shape == ref.shape and out.dtype == np.float32
    rel_log, rel_lin, _ = L.logmel_parity(ref, out)
    assert rel_log <= 1e-6 and rel_lin <= 1e-6, (rel_log, rel_lin)


def test_frame_count_contract():
    # T = 1 + N // hop : 47 975 samples -> 400 frames, 48 000 -> 401
    assert L.log_specgram(np.zeros(47975, np.float32)).shape == (400, 128)
    assert L.log_specgram(np.zeros(48000, np.float32)).shape == (401, 128)
    assert L.stft_sizes(24000, 10.0, 5.0) == (240, 120)
    assert L.stft_sizes(24000, 10.0, 2.5) == (240, 180)      # the EPIC-only quirk: hop = win - hop
    assert L.stft_sizes(24000, 10.0, 2.5, epic_quirk=False) == (240, 60)


def test_silence_hits_log_floor():
    out = L.log_specgram(np.zeros(4799, np.float32))
    assert np.allclose(out, np.log(np.float32(1e-6)), atol=1e-6)


def test_mel_filterbank_structure(gold):
    fb = L.mel_filterbank()
    assert fb.shape == (128, 1025) and fb.dtype == np.float32
    assert np.array_equal((fb > 0).sum(1), gold["mel_basis_nnz"])
    assert np.allclose(fb.sum(1), gold["mel_basis_rowsum"], rtol=1e-6)
    assert fb.min() >= 0 and fb.max() <= 1.0  # norm=None: unit-height triangles


def test_against_torchaudio():
    torchaudio = pytest.importorskip("torchaudio")
    ms = torchaudio.transforms.MelSpectrogram(
        sample_rate=24000, n_fft=2048, win_length=240, hop_length=120, f_min=0, f_max=12000, n_mels=128, power=1.0,
        center=True, pad_mode="constant", norm=None, mel_scale="htk")
    for name in ("noise0", "sine1k", "len5000"):
        w = _cases()[name]
        a = L.log_specgram(w)
        b = torch.log(ms(torch.from_numpy(w)) + 1e-6).T.numpy()
        rel_log, rel_lin, _ = L.logmel_parity(a, b)
        assert rel_log < 1e-4 and rel_lin < 1e-5, (name, rel_log, rel_lin)


def test_edge_padding():
    w = L.synthetic_waveforms(1, 5000, 97)[0]
    spec = L.extract_sound_feature(w, num_frames=400)
    assert spec.shape == (1, 400, 128)
    assert np.array_equal(spec[0, 41], spec[0, 399]) and not np.array_equal(spec[0, 40], spec[0, 41])


def test_slow_indices_match_golden(golden_dir):
    g = np.load(os.path.join(golden_dir, "pack_pathway.npz"))
    for key in g.files:
        T, a = key[1:].split("_a")
        assert np.array_equal(L.slow_indices(int(T), int(a)), g[key]), key
    idx = L.slow_indices(400, 4)
    assert len(idx) == 100 and idx[0] == 0 and idx[-1] == 399
    assert sorted(set(np.diff(idx))) == [4, 5] and (np.diff(idx) == 5).sum() == 3
