"""Weight checkpoints: a flat fp32 arena plus a shape table.

The reference has no persistence for tensor weights (plugins-Weights Weights.scala:39 keeps `data` as a plain var, and
TensorWeights.scala:139-153 only allocates); resuming a run or moving it to another GPU count needs one.  Replicas of
a data-parallel job are bit-identical (same counter-based init, SUM all-reduced gradients), so any rank's file restores
every rank of a job of any size.

File layout (little endian):
    "DLSW" u32 version(=1) u32 n_entries u32 reserved
    per entry: u32 name_len, name (utf-8), u32 rank, rank x u64 extent, u64 offset (floats, into the arena), u64 count
    padding to a 64-byte boundary, then the arena: every tensor's fp32 data, row-major, each starting on a 64-byte
    boundary (so the file can be mapped and uploaded without repacking)

`restore` writes INTO the existing weight and optimizer-state buffers (`Backend.write`), so captured steps that were
recorded on those buffers stay valid.
"""
from __future__ import annotations

import struct
from typing import Dict, List, Tuple

import numpy as np

MAGIC = b"DLSW"
VERSION = 1
_ALIGN_FLOATS = 16


def save_arrays(path: str, arrays: Dict[str, np.ndarray]) -> int:
    """Write named fp32 arrays; returns the number of bytes written."""
    table: List[Tuple[bytes, Tuple[int, ...], int, int]] = []
    offset = 0
    for name, a in arrays.items():
        a = np.asarray(a)
        if a.dtype != np.float32:
            raise ValueError("checkpoint entry %r is %s, expected float32" % (name, a.dtype))
        table.append((name.encode("utf-8"), tuple(int(s) for s in a.shape), offset, int(a.size)))
        offset += (a.size + _ALIGN_FLOATS - 1) // _ALIGN_FLOATS * _ALIGN_FLOATS
    head = bytearray(MAGIC + struct.pack("<III", VERSION, len(table), 0))
    for name, shape, off, count in table:
        head += struct.pack("<I", len(name)) + name + struct.pack("<I", len(shape))
        head += struct.pack("<%dQ" % len(shape), *shape) + struct.pack("<QQ", off, count)
    head += b"\0" * (-len(head) % 64)
    with open(path, "wb") as f:
        f.write(head)
        for (name, shape, off, count), a in zip(table, arrays.values()):
            flat = np.ascontiguousarray(a, dtype=np.float32).reshape(-1)
            f.write(flat.tobytes())
            f.write(b"\0" * (4 * (-count % _ALIGN_FLOATS)))
        return f.tell()


def load_arrays(path: str) -> Dict[str, np.ndarray]:
    with open(path, "rb") as f:
        blob = f.read()
    if len(blob) < 16 or blob[:4] != MAGIC:
        raise ValueError("%s is not a DLSW checkpoint" % path)
    version, n, _ = struct.unpack_from("<III", blob, 4)
    if version != VERSION:
        raise ValueError("checkpoint version %d, this build reads version %d" % (version, VERSION))
    pos = 16
    table = []
    try:
        for _ in range(n):
            (ln,) = struct.unpack_from("<I", blob, pos)
            pos += 4
            name = blob[pos:pos + ln].decode("utf-8")
            pos += ln
            (rank,) = struct.unpack_from("<I", blob, pos)
            pos += 4
            if rank > 64:
                raise ValueError("checkpoint entry %r has rank %d" % (name, rank))
            shape = struct.unpack_from("<%dQ" % rank, blob, pos)
            pos += 8 * rank
            off, count = struct.unpack_from("<QQ", blob, pos)
            pos += 16
            table.append((name, shape, off, count))
    except struct.error:
        raise ValueError("checkpoint %s: truncated shape table" % path) from None
    base = pos + (-pos % 64)
    out: Dict[str, np.ndarray] = {}
    for name, shape, off, count in table:
        if int(np.prod(shape, dtype=np.int64)) != count:
            raise ValueError("checkpoint entry %r: shape %r does not hold %d elements" % (name, shape, count))
        start = base + 4 * off
        if start + 4 * count > len(blob):
            raise ValueError("checkpoint %s: entry %r runs past the end of the file" % (path, name))
        out[name] = np.frombuffer(blob, dtype="<f4", count=count, offset=start).reshape(shape).copy()
    return out


# -- trainers ------------------------------------------------------------------------------------------------------------

def snapshot(trainer) -> Dict[str, np.ndarray]:
    """Weights in allocation order (`w0`, `w1`, ...), the stateful optimizer's buffers (`w0.s0`, ...) and its step count."""
    arrays: Dict[str, np.ndarray] = {}
    opt = trainer.P.optimizer
    for i, w in enumerate(trainer.weights):
        arrays["w%d" % i] = w.data.flat_array().reshape(w.data.shape)
        if opt is not None:
            for k, s in enumerate(opt.state_arrays(w)):
                arrays["w%d.s%d" % (i, k)] = s
    if opt is not None:
        # two 16-bit halves: each is exact in fp32, so the count survives up to 2^32 steps (a single fp32 stops at 2^24)
        arrays["optimizer.t"] = np.array([opt.t & 0xFFFF, (opt.t >> 16) & 0xFFFF], dtype=np.float32)
    return arrays


def save(path: str, trainer) -> int:
    return save_arrays(path, snapshot(trainer))


def restore(path: str, trainer) -> None:
    """Load `path` into the trainer's existing buffers; shapes must match exactly."""
    arrays = load_arrays(path)
    be = trainer.backend
    opt = trainer.P.optimizer
    for i, w in enumerate(trainer.weights):
        name = "w%d" % i
        if name not in arrays:
            raise ValueError("checkpoint has no entry %r (it holds %d weights)" % (name, sum(k[1:].isdigit() for k in arrays)))
        a = arrays[name]
        if list(a.shape) != list(w.data.shape):
            raise ValueError("checkpoint entry %r has shape %r, the model expects %r" % (name, list(a.shape), list(w.data.shape)))
        be.write(w.data._buffer(), a)
        if opt is not None:
            for k, s in enumerate(opt.state_of(w)):
                key = "w%d.s%d" % (i, k)
                if key not in arrays:
                    raise ValueError("checkpoint has no optimizer state %r" % key)
                be.write(s._buffer(), arrays[key])
    if opt is not None:
        if "optimizer.t" not in arrays:
            raise ValueError("checkpoint has no optimizer step count")
        t = arrays["optimizer.t"].reshape(-1)
        opt.t = int(t[0]) + (int(t[1]) << 16 if t.size > 1 else 0)      # one-element entries: files written before the split
    be.synchronize()
