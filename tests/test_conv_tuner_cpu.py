"""Host-side logic of the convolution autotuner (no GPU): candidate lists and tolerance of variants that do not fit."""
import pytest
import torch

import layerwise_b200  # noqa: F401
from layerwise_b200.ops import conv as C_


def test_candidate_lists():
    assert C_._candidates(64) == [64]
    assert C_._candidates(256) == [64, 128, 256]
    assert C_._candidates(192) == [64]
    # halo variant: one 64-channel chunk of K, stride 1, more than one tap that fits the tap table
    assert C_._halo_candidates(64, 128, 3, 1) == [1064, 1128]
    assert C_._halo_candidates(8, 64, 3, 1) == [1064]
    assert C_._halo_candidates(128, 128, 3, 1) == []
    assert C_._halo_candidates(64, 64, 1, 1) == []
    assert C_._halo_candidates(64, 64, 3, 2) == []
    assert C_._halo_candidates(64, 64, 5, 1) == []          # 25 taps > 16
    # two CTAs per SM: 64 and 128 wide; 64 only when a second tile is staged per sub-tile
    assert C_._two_candidates(256, False) == [2064, 2128]
    assert C_._two_candidates(256, True) == [2064]
    assert C_._two_candidates(64, False) == [2064]


def test_switches_remove_candidates(monkeypatch):
    monkeypatch.setitem(C_._HALO, "on", False)
    monkeypatch.setitem(C_._TWO, "on", False)
    assert C_._halo_candidates(64, 128, 3, 1) == [] and C_._two_candidates(256, False) == []


@pytest.mark.skipif(torch.cuda.is_available(), reason="the timing path needs no device only when tuning is off")
def test_tuner_is_inert_without_a_device(monkeypatch):
    calls = []
    monkeypatch.setitem(C_._TUNE_ON, "on", False)
    assert C_._tuned(("x", 1), 64, calls.append, [64, 128]) == 0 and calls == []


@pytest.mark.parametrize("span_w", [1, 2, 3, 4, 5])
def test_halo_staging_rows_are_a_permutation(span_w):
    """CPU restatement of the halo epilogue's row compaction (csrc/convx.cu, `dr`): accumulator row r = (r / 16, r % 16)
    of the 16-pixel-pitch box goes to row (r / 16) * Wb + r % 16 of the Wb-pixel-pitch staging tile when the column is
    an output pixel, and behind the patch otherwise -- every staging row is written exactly once."""
    PW, HB = 16, 8
    wb = PW - (span_w - 1)
    dest = []
    for r in range(128):
        r_w, r_h = r % PW, r // PW
        inside = r_w < wb and r_h < HB
        dest.append(r_h * wb + r_w if inside else wb * HB + r_h * (PW - wb) + (r_w - wb))
    assert sorted(dest) == list(range(128))
    assert all(d < wb * HB for r, d in enumerate(dest) if r % PW < wb)          # valid pixels fill the patch exactly


def test_tuning_table_round_trips_through_a_file(tmp_path, monkeypatch):
    monkeypatch.setattr(C_, "_TUNE", {("f", 224, 56, 56, 64, 256, 1, 1, 0, True): 2128, ("stem", 224, 112, 112, 64, True): 3064})
    path = str(tmp_path / "tune.json")
    assert C_.save_tuned(path) == 2
    monkeypatch.setattr(C_, "_TUNE", {("stem", 224, 112, 112, 64, True): 64})
    assert C_.load_tuned(path) == 1                                   # the entry already present wins
    assert C_._TUNE[("f", 224, 56, 56, 64, 256, 1, 1, 0, True)] == 2128 and C_._TUNE[("stem", 224, 112, 112, 64, True)] == 64
    # a tuned key short-circuits the timing loop
    assert C_._tuned(("f", 224, 56, 56, 64, 256, 1, 1, 0, True), 256, lambda h: (_ for _ in ()).throw(AssertionError("timed"))) == 2128
    (tmp_path / "bad.json").write_text('{"version": 7, "entries": []}')
    with pytest.raises(ValueError):
        C_.load_tuned(str(tmp_path / "bad.json"))
