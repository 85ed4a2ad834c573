"""Checkpoint container (TF V1 table format restated in sketch_rnn_b200/checkpoint.py): round trip,
checksums, key encoding, and the legacy .npz container.  No TensorFlow is available here, so the format
is only validated against itself (see the module docstring)."""
import io
import struct

import numpy as np
import pytest

from sketch_rnn_b200 import checkpoint as ck


def _vars(seed=0, big=False):
    rng = np.random.RandomState(seed)
    v = {"rnn_mdn/output_w": rng.randn(256 if big else 8, 147).astype(np.float32),
         "rnn_mdn/output_b": rng.randn(147).astype(np.float32),
         "rnn_mdn/MultiRNNCell/Cell0/BasicLSTMCell/Linear/Matrix": rng.randn(261 if big else 13, 1024 if big else 32).astype(np.float32),
         "rnn_mdn/MultiRNNCell/Cell0/BasicLSTMCell/Linear/Bias": np.zeros(1024 if big else 32, np.float32),
         "beta1_power": np.float32(0.81), "beta2_power": np.float32(0.998), "Variable": np.float32(0.005)}
    for k in list(v):
        if k.startswith("rnn_mdn"):
            v[k + "/Adam"] = np.zeros_like(v[k])
            v[k + "/Adam_1"] = np.abs(v[k]) * 1e-3
    return v


def test_crc32c_known_answers():
    assert ck.crc32c(b"") == 0
    assert ck.crc32c(b"123456789") == 0xE3069283        # standard CRC-32C check value
    assert ck.crc32c(b"\x00" * 32) == 0x8A9136AA         # RFC 3720 test vector
    assert ck.crc32c(b"56789", ck.crc32c(b"1234")) == 0xE3069283


@pytest.mark.parametrize("big", [False, True])
def test_round_trip(tmp_path, big):
    v = _vars(1, big)
    p = str(tmp_path / "model.ckpt-7")
    ck.write_checkpoint(p, v)
    blob = open(p, "rb").read()
    assert struct.unpack("<Q", blob[-8:])[0] == ck.TF_TABLE_MAGIC
    back = ck.read_checkpoint(p)
    assert set(back) == set(v)
    for k in v:
        assert back[k].dtype == np.float32 and back[k].shape == np.asarray(v[k]).shape, k
        assert np.array_equal(back[k], v[k]), k
    assert ck.list_checkpoint(p)["rnn_mdn/output_b"] == (147,)


def test_corruption_is_detected(tmp_path):
    p = str(tmp_path / "model.ckpt-1")
    ck.write_checkpoint(p, _vars(2))
    blob = bytearray(open(p, "rb").read())
    blob[100] ^= 0x40
    open(p, "wb").write(bytes(blob))
    with pytest.raises(IOError, match="CRC-32C"):
        ck.read_checkpoint(p)
    open(p, "wb").write(b"not a checkpoint at all, but long enough to have a footer.........")
    with pytest.raises(IOError, match="not a checkpoint"):
        ck.read_checkpoint(p)


def test_key_encoding_and_order():
    k = ck.encode_tensor_name_slice("a/b", 2)
    assert k == b"\x00" + b"a/b" + b"\x00\x01" + b"\x01\x02" + b"\x80\x7f\x80\x7f"
    assert ck.encode_tensor_name_slice("s", 0) == b"\x00s\x00\x01\x00"
    assert b"" < k  # the meta entry sorts first


def test_legacy_npz_container_still_loads(tmp_path):
    v = _vars(3)
    buf = io.BytesIO()
    np.savez(buf, **{k.replace("/", "|"): v[k] for k in v})
    p = str(tmp_path / "model.ckpt-0")
    open(p, "wb").write(buf.getvalue())
    back = ck.read_checkpoint(p)
    assert np.array_equal(back["rnn_mdn/output_w"], v["rnn_mdn/output_w"])
