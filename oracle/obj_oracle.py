"""Pure-Python restatement of `parseObj` (reference: src/Graphics/Ray/Geometry.hs:207-285).
TEST INFRASTRUCTURE ONLY: the checker for the library's C++ loader (rt_obj_parse) and the
generator of tests/golden/assets/*.npz.  Never imported by raytrace_b200/.

Returns (verts[n,3] float64, uvs[m,2] float64, faces[k,6] int32 = v0,t0,v1,t1,v2,t2 with t = -1
for Nothing) or raises ValueError(message) where the reference returns ``Left message``.
"""
from __future__ import annotations

import re
from typing import List, Optional, Tuple

import numpy as np

# Haskell's `read :: String -> Double` lexeme: digits [. digits] [e [+-] digits], optional leading '-'
_HS_DOUBLE = re.compile(r"^-?[0-9]+(\.[0-9]+)?([eE][-+]?[0-9]+)?$")
_WS = " \t\n\r\f\v"


def _read_double(tok: str) -> Optional[float]:
    return float(tok) if _HS_DOUBLE.match(tok) else None


def _words(s: str) -> List[str]:
    return s.split()  # Haskell `words`: split on isSpace


def _with_line(k: int, err: str) -> str:
    return f"line {k}: {err}"


def _extract_nat(s: str) -> Tuple[int, str]:
    n = 0
    while n < len(s) and s[n] in "0123456789":
        n += 1
    if n == 0:
        raise ValueError("expected number")
    return int(s[:n]), s[n:]


def _extract_int(s: str) -> Tuple[int, str]:
    if s.startswith("-"):
        i, rest = _extract_nat(s[1:])
        return -i, rest
    return _extract_nat(s)


def _process_ix(length: int, i: int) -> int:
    if 1 <= i <= length:
        return i - 1
    if -length <= i <= -1:
        return i + length
    raise ValueError(f"index out of bounds: {i}")


def _get_indices(num_vs: int, num_vts: int, tok: str) -> Tuple[int, int]:
    i, rest = _extract_int(tok)
    i2 = _process_ix(num_vs, i)
    if rest == "":
        return i2, -1
    if rest.startswith("//"):
        return i2, -1
    if rest.startswith("/"):
        j, _ = _extract_int(rest[1:])
        return i2, _process_ix(num_vts, j)
    raise ValueError(f"unexpected character '{rest[0]}'")


def parse_obj(text: str):
    lines = [ln.split("#", 1)[0] for ln in text.split("\n")]   # removeComments . lines
    if text.endswith("\n"):
        lines = lines[:-1]                                      # `lines` drops the final empty piece
    v_lines, vt_lines, f_lines = [], [], []
    for k, ln in enumerate(lines, start=1):
        if ln.startswith("v "):
            v_lines.append((k, ln[2:]))
        elif ln.startswith("vt "):
            vt_lines.append((k, ln[3:]))
        elif ln.startswith("f "):
            f_lines.append((k, ln[2:]))
    verts = []
    for k, ln in v_lines:
        w = _words(ln)
        vals = [_read_double(t) for t in w[:3]]
        if len(vals) < 3 or any(x is None for x in vals):
            raise ValueError(_with_line(k, "invalid 'v' statement"))
        verts.append(vals)
    uvs = []
    for k, ln in vt_lines:
        w = _words(ln)
        if len(w) == 1 and _read_double(w[0]) is not None:
            uvs.append([_read_double(w[0]), 0.0])
        elif len(w) >= 2 and _read_double(w[0]) is not None and _read_double(w[1]) is not None:
            uvs.append([_read_double(w[0]), _read_double(w[1])])
        else:
            raise ValueError(_with_line(k, "invalid 'vt' statement"))
    faces = []
    for k, ln in f_lines:
        try:
            idx = [_get_indices(len(verts), len(uvs), t) for t in _words(ln)]
        except ValueError as e:
            raise ValueError(_with_line(k, str(e))) from None
        if len(idx) < 3:
            raise ValueError(_with_line(k, "invalid 'f' statement (fewer than 3 vertices)"))
        i0, rest = idx[0], idx[1:]
        for a, b in zip(rest, rest[1:]):                        # fan triangulation (`pairs`)
            faces.append([i0[0], i0[1], a[0], a[1], b[0], b[1]])
    return (np.array(verts, dtype=np.float64).reshape(-1, 3),
            np.array(uvs, dtype=np.float64).reshape(-1, 2),
            np.array(faces, dtype=np.int32).reshape(-1, 6))
