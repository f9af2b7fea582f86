"""Input/output helpers for the matFact file format (host side, Python mirror).

The format is the reference's (matFact.c:72-105): whitespace separated tokens
``iters alpha features users items nnz`` followed by ``nnz`` triples
``row col value``.  The reference's parser is token based (``fscanf``), so line
breaks carry no meaning; this reader is token based too.  Entries keep the
file order and the reference's 16-byte ``non_zero_entry`` layout
(datatypes.h:10-15) so the array can be handed to the C ABI as is.
"""
from __future__ import annotations

import gzip
import io
from dataclasses import dataclass

import numpy as np

ENTRY_DTYPE = np.dtype([("row", "<i4"), ("col", "<i4"), ("value", "<f8")], align=True)
assert ENTRY_DTYPE.itemsize == 16


@dataclass
class Instance:
    iters: int
    alpha: float
    features: int
    users: int
    items: int
    entries: np.ndarray  # ENTRY_DTYPE, file order

    @property
    def nnz(self) -> int:
        return int(self.entries.shape[0])


def _read_text(path) -> str:
    path = str(path)
    if path.endswith(".gz"):
        with gzip.open(path, "rt") as fh:
            return fh.read()
    with open(path, "r") as fh:
        return fh.read()


def parse_instance(text: str) -> Instance:
    tok = text.split()
    if len(tok) < 6:
        raise ValueError("truncated header")
    iters = int(tok[0])
    alpha = float(tok[1])
    features = int(tok[2])
    users, items, nnz = int(tok[3]), int(tok[4]), int(tok[5])
    body = tok[6:6 + 3 * nnz]
    if len(body) != 3 * nnz:
        raise ValueError("Error in non-zero entry.")
    ent = np.zeros(nnz, dtype=ENTRY_DTYPE)
    if nnz:
        ent["row"] = np.array(body[0::3], dtype=np.int64).astype(np.int32)
        ent["col"] = np.array(body[1::3], dtype=np.int64).astype(np.int32)
        ent["value"] = np.array(body[2::3], dtype=np.float64)
    return Instance(iters, alpha, features, users, items, ent)


def read_instance(path) -> Instance:
    return parse_instance(_read_text(path))


def write_instance(path, inst: Instance, value_fmt: str = "%.1f") -> None:
    """Write an instance in the reference's layout (one token group per line)."""
    buf = io.StringIO()
    buf.write("%d\n%s\n%d\n%d %d %d\n" % (inst.iters, repr(float(inst.alpha)), inst.features,
                                         inst.users, inst.items, inst.nnz))
    e = inst.entries
    vals = [value_fmt % v for v in e["value"]]
    buf.write("\n".join("%d %d %s" % t for t in zip(e["row"].tolist(), e["col"].tolist(), vals)))
    buf.write("\n")
    data = buf.getvalue()
    path = str(path)
    if path.endswith(".gz"):
        with gzip.open(path, "wt") as fh:
            fh.write(data)
    else:
        with open(path, "w") as fh:
            fh.write(data)


def format_output(best) -> str:
    """stdout of the reference for a recommendation vector (matFact.c:24-25):
    one ``%d\\n`` per user whose entry is not -1."""
    return "".join("%d\n" % b for b in np.asarray(best).tolist() if b != -1)


def read_golden_out(path) -> str:
    """A samples/X.out file with blank lines dropped (one upstream golden,
    inst200-10000-50-100-300.out, carries a trailing blank line)."""
    txt = _read_text(path)
    return "".join(line.strip() + "\n" for line in txt.splitlines() if line.strip())
