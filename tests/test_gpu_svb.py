"""pgm_upload_svb on the device: the svb-zd streams decode to exactly the int16 samples slow5lib's decoder gives
(bit-exact, integer work), and a whole gmove pass over them equals the pass over plain raw[]."""
import numpy as np
import pytest

import oracle_lib as ol
from cases import parity_cases
from poregen_b200 import _capi as capi, synth
from poregen_b200.batch import GmoveOptions, ReadBatch
from poregen_b200.gmove import GmoveEngine, PgmError

pytestmark = pytest.mark.gpu


_aligned = synth.align_raw


def _check_decode(eng, batch):
    blob, off = synth.svb_streams(batch)
    eng.upload_svb(batch, blob, off)
    got = eng.device_raw(batch.raw.size)
    for r in batch.reads:                                   # gaps between records are not written: compare per record
        o, n = int(r["raw_off"]), int(r["raw_len"])
        assert np.array_equal(got[o:o + n], batch.raw[o:o + n])
    return blob, off


@pytest.mark.parametrize("align", [1, 8])
def test_decode_matches_samples(built, align):
    cfg = synth.SynthConfig(n_reads=300, median_samples=6000, seed=41, spike_frac=0.001)
    batch = synth.make_batch(cfg)
    if align > 1:
        batch = _aligned(batch, align)
    opts = GmoveOptions(kmer_size=5, sample_limit=20, file_limit=1024, index_end=1024).resolve_range()
    with GmoveEngine(opts, device=0) as eng:
        blob, off = _check_decode(eng, batch)
    # and the streams are what slow5lib's decoder reads (oracle restatement of it)
    for i in (0, 1, batch.n_reads - 1):
        s = blob[int(off[i]):int(off[i + 1])].tobytes()
        o, n = int(batch.reads[i]["raw_off"]), int(batch.reads[i]["raw_len"])
        assert np.array_equal(ol.svb_zd_decode(s, n), batch.raw[o:o + n])


def test_decode_edge_lengths_and_wide_values(built):
    rng = np.random.default_rng(9)
    lens = [0, 1, 2, 3, 4, 5, 7, 8, 9, 31, 32, 33, 255, 256, 257, 511, 512, 513, 1023, 1025, 4099]
    sigs = [rng.integers(-300, 300, size=n).astype(np.int16) for n in lens]
    sigs.append(rng.integers(-32768, 32768, size=5000).astype(np.int16))                 # 2- and 3-byte values, chunks of ~600 bytes
    sigs.append(np.tile(np.array([32767, -32768], np.int16), 1500))                      # every value 3 bytes: the widest chunk
    sigs.append(np.full(3000, -7, np.int16))
    reads = np.zeros(len(sigs), dtype=np.dtype(capi.READ_DTYPE))
    pos = 0
    for i, s in enumerate(sigs):
        reads[i]["raw_off"], reads[i]["raw_len"] = pos, s.size
        reads[i]["digitisation"], reads[i]["range"], reads[i]["stride"], reads[i]["fastq_len"] = 8192.0, 1467.61, 5, 100
        pos += s.size
    batch = ReadBatch(capi.PGM_MODE_MOVES_CHAR, np.concatenate(sigs), np.zeros(0, np.uint8), np.zeros(0, np.uint8), reads,
                      [f"r{i}" for i in range(len(sigs))])
    opts = GmoveOptions(kmer_size=5, sample_limit=20, file_limit=1024, index_end=1024).resolve_range()
    with GmoveEngine(opts, device=0) as eng:
        _check_decode(eng, batch)
        _check_decode(eng, _aligned(batch, 8))


def test_malformed_stream_fails_loudly(built):
    cfg = synth.SynthConfig(n_reads=8, median_samples=3000, seed=42)
    batch = synth.make_batch(cfg)
    blob, off = synth.svb_streams(batch)
    opts = GmoveOptions(kmer_size=5, sample_limit=20, file_limit=1024, index_end=1024).resolve_range()
    with GmoveEngine(opts, device=0) as eng:
        bad = blob.copy()
        bad[int(off[3])] ^= 1                                # record 3 announces another count
        eng.upload_svb(batch, bad, off)
        with pytest.raises(PgmError) as e:
            eng.device_raw(batch.raw.size)
        assert e.value.code == capi.PGM_ERR_INPUT
        off2 = off.copy()
        off2[4:] += np.uint64(1)                             # record 3's stream gets one byte too many
        bad2 = np.concatenate([blob[:int(off[4])], np.zeros(1, np.uint8), blob[int(off[4]):]])
        eng.upload_svb(batch, bad2, off2)
        eng.count(); eng.emit()
        with pytest.raises(PgmError) as e:
            eng.download()
        assert e.value.code == capi.PGM_ERR_INPUT
        eng.upload_svb(batch, blob, off)                     # and the context recovers
        eng.count(); eng.emit(); eng.download()


CASES = [c for c in parity_cases() if c[0] in ("tsv_k5_sc1", "sam_k5_sc1", "ss_rna_k5_sc1", "ss_dna_k6_sc0", "tsv_offsets")]


@pytest.mark.parametrize("label,cfg,opts", CASES, ids=[c[0] for c in CASES])
def test_gmove_over_svb_equals_oracle(built, label, cfg, opts):
    batch = synth.make_batch(cfg)
    ref = ol.oracle_run(opts, batch)
    blob, off = synth.svb_streams(batch)
    with GmoveEngine(opts, device=0) as eng:
        eng.upload_svb(batch, blob, off)
        eng.count(); eng.emit()
        res = eng.download()
    assert np.array_equal(res.slot_total, ref.slot_total)
    assert np.array_equal(res.ev_len, ref.ev_len) and np.array_equal(res.ev_start, ref.ev_start) and np.array_equal(res.ev_read, ref.ev_read)
    assert np.array_equal(res.samples.view(np.uint64), ref.samples.view(np.uint64))
    assert np.array_equal(res.read_flags, ref.read_flags)
