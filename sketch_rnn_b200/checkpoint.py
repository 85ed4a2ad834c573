"""Checkpoint files under save/<dataset>/ (reference train.py:72-75, sample.py:59-64).

`write_checkpoint` / `read_checkpoint` move a {tf_variable_name: array} dict (trainables, Adam slots,
beta powers, learning rate -- everything tf.train.Saver(tf.all_variables()) would hold) to and from
`model.ckpt-<step>`.  The container written here is a NumPy .npz stream stored under the reference's
file name; `read_checkpoint` dispatches on the file magic so that a reader for TensorFlow's V1 table
format (the shipped save/kanji/model.ckpt-0, absent from the reference mount) can be added beside it.
"""
import io
import zipfile

import numpy as np

TF_TABLE_MAGIC = 0xdb4775248b80fb57  # footer of a TF V1 (SSTable) checkpoint


def write_checkpoint(path, variables):
    buf = io.BytesIO()
    np.savez(buf, **{k.replace("/", "|"): np.asarray(v) for k, v in variables.items()})
    with open(path, "wb") as f:
        f.write(buf.getvalue())
    return path


def read_checkpoint(path):
    with open(path, "rb") as f:
        blob = f.read()
    if blob[:2] == b"PK":
        with np.load(io.BytesIO(blob)) as z:
            return {k.replace("|", "/"): z[k] for k in z.files}
    if len(blob) >= 8 and int.from_bytes(blob[-8:], "little") == TF_TABLE_MAGIC:
        raise NotImplementedError("%s is a TensorFlow V1 table checkpoint; its reader is not built yet "
                                  "(SURVEY.md 8f-1)" % path)
    raise IOError("%s is not a checkpoint this library can read" % path)
