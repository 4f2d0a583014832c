"""Synthetic 3DGS scenes: deterministic, counter-based, chunked.

The reference ships no scene and no generator (its only model path is an absolute
path on the author's machine, src/main.mm:69-70), so the benchmark family is
defined here.  Files are standard 3DGS binary-little-endian PLYs with the property
names the reference loader matches (src/ply_loader.cpp:59-81).

Every value is a pure function of (seed, vertex index, slot): splitmix64 of a
counter, so any chunk of any scene can be regenerated independently.
"""
from __future__ import annotations

import numpy as np

SLOTS = 128  # counter stride per vertex (>= 2 * number of random draws)
_GOLD = np.uint64(0x9E3779B97F4A7C15)

PROPS_FULL = (
    ["x", "y", "z", "nx", "ny", "nz", "f_dc_0", "f_dc_1", "f_dc_2"]
    + [f"f_rest_{i}" for i in range(45)]
    + ["opacity", "scale_0", "scale_1", "scale_2", "rot_0", "rot_1", "rot_2", "rot_3"]
)
PROPS_NOSH = [p for p in PROPS_FULL if not p.startswith("f_rest_")]

CONFIG_SEEDS = {1: 0x5EED0001, 2: 0x5EED0002, 3: 0x5EED0003, 4: 0x5EED0004, 5: 0x5EED0005}


def _mix(z: np.ndarray) -> np.ndarray:
    with np.errstate(over="ignore"):
        z = (z ^ (z >> np.uint64(30))) * np.uint64(0xBF58476D1CE4E5B9)
        z = (z ^ (z >> np.uint64(27))) * np.uint64(0x94D049BB133111EB)
        return z ^ (z >> np.uint64(31))


def _uniform(seed: int, index: np.ndarray, slot: int) -> np.ndarray:
    """float64 in (0, 1), keyed by (seed, index, slot)."""
    with np.errstate(over="ignore"):
        ctr = index.astype(np.uint64) * np.uint64(SLOTS) + np.uint64(slot + 1)
        z = _mix(np.uint64(seed) + ctr * _GOLD)
    return ((z >> np.uint64(11)).astype(np.float64) + 0.5) * (1.0 / 9007199254740992.0)


def _normal(seed: int, index: np.ndarray, slot: int) -> np.ndarray:
    """Box-Muller from slots (2*slot, 2*slot+1)."""
    u1 = _uniform(seed, index, 2 * slot)
    u2 = _uniform(seed, index, 2 * slot + 1)
    return np.sqrt(-2.0 * np.log(u1)) * np.cos(2.0 * np.pi * u2)


def blob_centres(seed: int, n_blobs: int = 64) -> np.ndarray:
    idx = np.arange(n_blobs, dtype=np.uint64) + np.uint64(1 << 40)
    return np.stack([(_uniform(seed, idx, s) * 2.0 - 1.0) * 3.5 for s in range(3)], axis=1)


def gen_chunk(n_total: int, seed: int, start: int, count: int, sh_rest: bool = True) -> np.ndarray:
    """Rows [start, start+count) of the raw PLY table (float32, PLY property order)."""
    idx = np.arange(start, start + count, dtype=np.uint64)
    centres = blob_centres(seed)
    in_blob = _uniform(seed, idx, 0) < 0.8
    which = np.minimum((_uniform(seed, idx, 1) * 64.0).astype(np.int64), 63)
    pos = np.empty((count, 3))
    for a in range(3):
        blob = centres[which, a] + 0.35 * _normal(seed, idx, 2 + a)      # slots 4..9
        unif = (_uniform(seed, idx, 10 + a) * 2.0 - 1.0) * 4.5
        pos[:, a] = np.where(in_blob, blob, unif)
    pos = np.clip(pos, -4.9, 4.9)                                        # D12: crop is a no-op
    s0 = 0.6 * float(n_total) ** (-1.0 / 3.0)
    cols = {
        "x": pos[:, 0], "y": pos[:, 1], "z": pos[:, 2],
        "nx": np.zeros(count), "ny": np.zeros(count), "nz": np.zeros(count),
    }
    for a in range(3):
        cols[f"f_dc_{a}"] = _normal(seed, idx, 8 + a)                    # slots 16..21
        cols[f"scale_{a}"] = np.log(s0) + 0.5 * _normal(seed, idx, 11 + a)
    cols["opacity"] = 1.5 * _normal(seed, idx, 14)
    for a in range(4):
        cols[f"rot_{a}"] = _normal(seed, idx, 15 + a)
    props = PROPS_FULL if sh_rest else PROPS_NOSH
    if sh_rest:
        for k in range(45):
            cols[f"f_rest_{k}"] = 0.1 * _normal(seed, idx, 19 + k)       # slots 38..127
    return np.stack([cols[p] for p in props], axis=1).astype("<f4")


def write_ply(path: str, n: int, seed: int, sh_rest: bool = True, chunk: int = 1 << 18) -> None:
    props = PROPS_FULL if sh_rest else PROPS_NOSH
    header = "ply\nformat binary_little_endian 1.0\nelement vertex %d\n" % n
    header += "".join(f"property float {p}\n" for p in props) + "end_header\n"
    with open(path, "wb") as f:
        f.write(header.encode("ascii"))
        for s in range(0, n, chunk):
            f.write(gen_chunk(n, seed, s, min(chunk, n - s), sh_rest).tobytes())


def raw_to_instances(raw: np.ndarray, sh_rest: bool = True):
    """Apply the loader's activations (src/ply_loader.cpp:116-139) and the host
    constructor's field copy + crop (src/instanced_splat_renderer.mm:360-388) in
    numpy.  Used to build big benchmark scenes without a 1.5 GB file; the C++
    PLYLoader is the product path and tests check that both agree.
    Returns (instances [m,14] f32, sh_rest [m,45] f32 or None)."""
    props = PROPS_FULL if sh_rest else PROPS_NOSH
    c = {p: raw[:, i].astype(np.float32) for i, p in enumerate(props)}
    one = np.float32(1.0)
    opacity = one / (one + np.exp(-c["opacity"], dtype=np.float32))
    scale = [np.exp(c[f"scale_{a}"], dtype=np.float32) for a in range(3)]
    dc = np.stack([c[f"f_dc_{a}"] for a in range(3)], axis=1)
    nonzero = np.any(dc != 0, axis=1, keepdims=True)
    rgb = np.clip(np.float32(0.5) + np.float32(0.28209479177387814) * dc, 0, 1).astype(np.float32)
    rgb = np.where(nonzero, rgb, dc)
    inst = np.stack(
        [c["rot_0"], c["rot_1"], c["rot_2"], c["rot_3"], scale[0], scale[1], scale[2],
         c["x"], c["y"], c["z"], rgb[:, 0], rgb[:, 1], rgb[:, 2], opacity], axis=1).astype(np.float32)
    keep = (np.abs(c["x"]) < 5) & (np.abs(c["y"]) < 5) & (np.abs(c["z"]) < 5)
    sh = None
    if sh_rest:
        sh = np.stack([c[f"f_rest_{k}"] for k in range(45)], axis=1)[keep]
    return np.ascontiguousarray(inst[keep]), sh


def make_instances(n: int, seed: int, sh_rest: bool = False, chunk: int = 1 << 18):
    """Whole scene as a SplatInstance array (56-byte rows), generated in chunks."""
    parts, shs = [], []
    for s in range(0, n, chunk):
        inst, sh = raw_to_instances(gen_chunk(n, seed, s, min(chunk, n - s), sh_rest), sh_rest)
        parts.append(inst)
        if sh is not None:
            shs.append(sh)
    return np.concatenate(parts), (np.concatenate(shs) if shs else None)
