"""svb-zd signal streams: the oracle's decoder and the numpy encoder that makes the test and bench
inputs, pinned against slow5lib's own compressor / decompressor (oracle/_ref/libsvbref.so, compiled from the reference's
submodule by oracle/build_ref.sh) and against known-answer streams worked out from the format by hand."""
import numpy as np
import pytest

import oracle_lib as ol
from poregen_b200 import synth


def _signals():
    rng = np.random.default_rng(5)
    walk = np.cumsum(rng.integers(-40, 41, size=5000)).astype(np.int16)
    yield "walk", walk
    yield "empty", np.zeros(0, np.int16)
    for n in (1, 2, 3, 4, 5, 7, 8, 9, 255, 256, 257, 511, 513):
        yield f"n{n}", rng.integers(-300, 300, size=n).astype(np.int16)
    yield "extremes", np.array([32767, -32768, 32767, -32768, 0, -1, 1, 127, -128, 128, 16383, -16384], np.int16)
    yield "full_range", rng.integers(-32768, 32768, size=3000).astype(np.int16)     # 2- and 3-byte values
    yield "constant", np.full(1000, 517, np.int16)                                   # all-zero differences: 1 byte each
    yield "nanopore_like", (np.repeat(rng.normal(500, 80, 400), 12)[:4700] + rng.normal(0, 9, 4700)).astype(np.int16)


SIGNALS = list(_signals())


def test_known_answer_streams():
    # count=3 | key byte: codes (0,1,0) -> 0b00_00_01_00 = 0x04 | data: 0x02 (zz 2 -> +1), 0x2C 0x01 (zz 300 -> +150), 0x01 (zz 1 -> -1)
    s = bytes([3, 0, 0, 0, 0x04, 0x02, 0x2C, 0x01, 0x01])
    assert ol.svb_zd_decode(s, 3).tolist() == [1, 151, 150]
    assert synth.svb_zd_encode(np.array([1, 151, 150], np.int16)) == s
    # a 3-byte value: difference -40000 (int32) -> zz 79999 = 0x01387F, truncated to int16 on store like the reference
    x = np.array([20000, -20000], np.int16)
    enc = synth.svb_zd_encode(x)
    assert enc == bytes([2, 0, 0, 0, 0x09, 0x40, 0x9C, 0x7F, 0x38, 0x01])
    assert ol.svb_zd_decode(enc, 2).tolist() == [20000, -20000]


@pytest.mark.parametrize("label,x", SIGNALS, ids=[s[0] for s in SIGNALS])
def test_roundtrip_and_reference(label, x):
    enc = synth.svb_zd_encode(x)
    dec = ol.svb_zd_decode(enc, x.size)
    assert dec is not None and np.array_equal(dec, x)
    if ol.svbref() is None:
        pytest.skip("oracle/_ref/libsvbref.so not built (no /root/reference here)")
    assert enc == ol.svbref_compress(x), "numpy encoder differs from slow5lib's ptr_compress_svb_zd"
    ref = ol.svbref_depress(enc)
    assert ref is not None and np.array_equal(ref, x)


def test_malformed_streams_are_rejected_like_slow5lib():
    x = np.arange(100, dtype=np.int16) * 7
    enc = synth.svb_zd_encode(x)
    bad = [enc[:-1], enc + b"\0", enc[:3], bytes([101, 0, 0, 0]) + enc[4:]]
    for s in bad:
        assert ol.svb_zd_decode(s, 100) is None
    assert ol.svb_zd_decode(enc, 99) is None          # the caller's length disagrees with the stream's
    if ol.svbref() is not None:
        assert ol.svbref_depress(enc[:-1]) is None and ol.svbref_depress(enc + b"\0") is None


def test_reference_reads_the_compressed_blow5_the_writer_makes(tmp_path):
    """The test/bench BLOW5 writer with signal_press svb-zd (+ zlib records) produces files the UNMODIFIED reference
    (slow5lib inside oracle/_ref/poregen_ref) reads to the same dump as the uncompressed file."""
    import os
    from cases import parity_cases
    if not os.path.exists(ol.REF_BIN):
        pytest.skip("compiled reference (oracle/_ref) not present")
    _, cfg, opts = next(c for c in parity_cases() if c[0] == "tsv_k5_sc1")
    cfg.n_reads = 12
    batch = synth.make_batch(cfg)
    outs = []
    for i, (sp, rp) in enumerate((("none", "none"), ("svb-zd", "none"), ("svb-zd", "zlib"))):
        d = tmp_path / f"v{i}"
        d.mkdir()
        synth.write_blow5(batch, str(d / "reads.blow5"), signal_press=sp, record_press=rp)
        fl = synth.write_alignment(batch, str(d))
        outs.append(ol.run_reference(opts, str(d / "reads.blow5"), fl["align"], str(d / "ref"), fastq=fl["fastq"]))
    assert outs[0] == outs[1] == outs[2] and len(outs[0]) > 1
