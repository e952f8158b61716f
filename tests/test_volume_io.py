"""Volume files: PVM / DDS / RAW reader and 16 -> 8 bit quantiser of libvr5.so (host code, no GPU)
against the reference's own loader (where /root/reference exists) and against fixtures that loader produced."""
import gzip
import hashlib
import json
import os

import numpy as np
import pytest

import oracle

HERE = os.path.dirname(os.path.abspath(__file__))
BUCKY = "/root/reference/VolumeRendering/Bucky.pvm"
FIX = os.path.join(HERE, "golden", "volume_io.json")


def _synth16(w, h, d, seed):
    """Smooth 16-bit field + noise, big-endian bytes as a PVM with 2 components stores them."""
    rng = np.random.default_rng(seed)
    z, y, x = np.meshgrid(np.arange(d), np.arange(h), np.arange(w), indexing="ij")
    v = (3000 + 2500 * np.sin(x / 3.0) * np.cos(y / 4.0) + 900 * np.sin(z / 2.5) + rng.integers(0, 200, size=(d, h, w))).astype(np.int64)
    v = np.clip(v, 0, 65535).astype(">u2")
    return v.tobytes()


@pytest.mark.skipif(not os.path.exists(BUCKY), reason="needs the reference's Bucky.pvm")
def test_bucky_pvm_decodes_to_the_reference_bytes(R, bucky, ref):
    vox, scale = R.load_pvm(BUCKY)
    assert vox.shape == (32, 32, 32, 1) and scale == (1.0, 1.0, 1.0)
    assert np.array_equal(vox[..., 0], bucky)                      # fixture decoded by the reference (tests/golden)
    assert np.array_equal(vox, ref.read_pvm(BUCKY))                # the reference's readPVMvolume, live
    assert np.array_equal(R.load_volume(BUCKY), bucky)
    assert int(vox.sum()) == 1174726


def test_dds_round_trip_all_modes(R):
    rng = np.random.default_rng(3)
    smooth = (128 + 100 * np.sin(np.arange(40000) / 37.0)).astype(np.uint8).tobytes()
    noisy = rng.integers(0, 256, size=10007, dtype=np.uint8).tobytes()
    for data in (smooth, noisy, b"", b"\x07", bytes(range(256)) * 3):
        for skip in (1, 2, 3, 4):
            for strip in (1, 2, 64, 4097):
                for v3e in (False, True):
                    stream = R.dds_encode(data, skip, strip, v3e)
                    assert stream[:8] == (b"DDS v3e\n" if v3e else b"DDS v3d\n")
                    assert R.dds_decode(stream) == data, (len(data), skip, strip, v3e)
    assert len(R.dds_encode(smooth, 1, 200)) < len(smooth) // 2    # the predictor earns its keep on smooth data


@pytest.mark.skipif(not oracle.ref_available(), reason="needs the reference build")
def test_our_streams_decode_with_the_reference_decoder(R, ref, tmp_path):
    """Encoder -> the reference's readPVMvolume: the stream layout is the reference's, not a private dialect."""
    rng = np.random.default_rng(9)
    vol = rng.integers(0, 256, size=(7, 9, 11), dtype=np.uint8)
    vol[2:5] = (np.arange(11) * 9).astype(np.uint8)
    for i, (skip, strip, v3e, header) in enumerate([(1, 11, False, b"PVM\n11 9 7\n1\n"), (2, 99, True, b"PVM2\n11 9 7\n1 1 2\n1\n"), (4, 1, False, b"PVM\n# comment\n11 9 7\n1\n")]):
        path = str(tmp_path / ("v%d.pvm" % i))
        with open(path, "wb") as f:
            f.write(R.dds_encode(header + vol.tobytes(), skip, strip, v3e))
        assert np.array_equal(ref.read_pvm(path)[..., 0], vol)
        mine, scale = R.load_pvm(path)
        assert np.array_equal(mine[..., 0], vol)
        assert scale == ((1.0, 1.0, 2.0) if i == 1 else (1.0, 1.0, 1.0))


def test_pvm3_and_raw_and_errors(R, tmp_path):
    vol = np.arange(2 * 3 * 4, dtype=np.uint8).reshape(2, 3, 4)
    p3 = str(tmp_path / "v3.pvm")
    with open(p3, "wb") as f:
        f.write(b"PVM3\n4 3 2\n1 2 3\n1\n" + vol.tobytes() + b"desc\0court\0param\0comment\0")
    got, scale = R.load_pvm(p3)
    assert np.array_equal(got[..., 0], vol) and scale == (1.0, 2.0, 3.0)
    raw = str(tmp_path / "v.raw")
    vol.tofile(raw)
    assert np.array_equal(R.load_raw(raw, 4, 3, 2)[..., 0], vol)
    with pytest.raises(R.Vr5Error):
        R.load_raw(raw, 4, 3, 3)                                   # ModelBase.cpp:77-82: size mismatch is an error
    with pytest.raises(R.Vr5Error):
        R.load_pvm(str(tmp_path / "missing.pvm"))                  # ModelBase.cpp:51-54: 1, no exit
    bad = str(tmp_path / "bad.pvm")
    with open(bad, "wb") as f:
        f.write(b"PVM\n4 3 2\n1\n" + vol.tobytes()[:-1])
    with pytest.raises(R.Vr5Error):
        R.load_pvm(bad)
    # header fields whose product wraps modulo 2^64 (2^22 * 2^22 * 2^20 * 1 = 2^64 -> 0 voxels) must not pass
    # the size check with a tiny buffer; neither must dimensions a ushort3 cannot hold
    for hdr in (b"PVM\n4194304 4194304 1048576\n1\n", b"PVM\n65536 1 1\n1\n", b"PVM\n2 2 2\n5\n"):
        huge = str(tmp_path / "huge.pvm")
        with open(huge, "wb") as f:
            f.write(hdr + bytes(64))
        with pytest.raises(R.Vr5Error, match="65535"):
            R.load_pvm(huge)
    three = str(tmp_path / "three.pvm")
    with open(three, "wb") as f:
        f.write(b"PVM\n2 2 2\n3\n" + bytes(24))
    with pytest.raises(R.Vr5Error):
        R.load_volume(three)                                       # ModelBase.cpp:85-89: more than two components


def test_quantiser_matches_fixtures_and_reference(R):
    fixtures = json.load(open(FIX))
    for case in fixtures["quantize"]:
        w, h, d = case["dims"]
        data = _synth16(w, h, d, case["seed"])
        for linear in (False, True):
            q = R.quantize16(np.frombuffer(data, dtype=np.uint8), w, h, d, linear)
            assert hashlib.sha256(q.tobytes()).hexdigest() == case["linear" if linear else "nonlinear"], (case["dims"], linear)
            if oracle.ref_available():
                assert np.array_equal(q, oracle.Ref().quantize16(np.frombuffer(data, dtype=np.uint8), w, h, d, linear))


@pytest.mark.skipif(not oracle.ref_available(), reason="needs the reference build")
def test_two_component_pvm_is_quantised_like_load_model(R, ref, tmp_path):
    w, h, d = 12, 10, 8
    data = _synth16(w, h, d, 4)
    path = str(tmp_path / "v16.pvm")
    with open(path, "wb") as f:
        f.write(R.dds_encode(b"PVM\n12 10 8\n2\n" + data, 2, 12))
    assert ref.load_model(path) == 0
    assert np.array_equal(R.load_volume(path), ref.volume())
