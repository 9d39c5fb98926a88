"""Packed record layouts — Python mirror of ``include/garb200_layout.h``.

The C header is authoritative; ``tests/test_layout.py`` checks this mirror against the
offsets reported by the C-ABI library (``garb200_layouts``) and by the oracle.

One knot of the reference (``LqrKnotTpl``, /root/reference/include/aligator/gar/
lqr-problem.hpp:32-56, 17 separately allocated matrices) becomes ONE contiguous record of
doubles so a single bulk copy stages it into shared memory.
"""
from __future__ import annotations

from dataclasses import dataclass

import numpy as np


def _r2(x: int) -> int:
    return (x + 1) & ~1


KNOT_FIELDS = ("Q", "S", "R", "q", "r", "A", "B", "E", "f", "C", "D", "d",
               "Gth", "Gx", "Gu", "gamma")


@dataclass(frozen=True)
class KnotLayout:
    nx: int
    nu: int
    nc: int
    nth: int

    def shapes(self, nu=None, nc=None):
        """(rows, cols) of every field for a knot with its own nu / nc (<= layout's)."""
        nx, nth = self.nx, self.nth
        nu = self.nu if nu is None else nu
        nc = self.nc if nc is None else nc
        return {
            "Q": (nx, nx), "S": (nx, nu), "R": (nu, nu), "q": (nx, 1), "r": (nu, 1),
            "A": (nx, nx), "B": (nx, nu), "E": (nx, nx), "f": (nx, 1),
            "C": (nc, nx), "D": (nc, nu), "d": (nc, 1),
            "Gth": (nth, nth), "Gx": (nx, nth), "Gu": (nu, nth), "gamma": (nth, 1),
        }

    @property
    def offsets(self):
        o = 0
        out = {}
        for name, (r, c) in self.shapes().items():
            out[name] = o
            o = _r2(o + r * c)
        out["stride"] = o
        return out

    @property
    def stride(self) -> int:
        return self.offsets["stride"]

    def pack(self, rec: np.ndarray, fields: dict, nu=None, nc=None) -> None:
        """Write column-major blocks of one knot into ``rec`` (1-D, len >= stride)."""
        off = self.offsets
        for name, (r, c) in self.shapes(nu, nc).items():
            if name not in fields or r * c == 0:
                continue
            a = np.asarray(fields[name], dtype=np.float64).reshape(r, c)
            rec[off[name]: off[name] + r * c] = a.reshape(-1, order="F")

    def unpack(self, rec: np.ndarray, nu=None, nc=None) -> dict:
        off = self.offsets
        out = {}
        for name, (r, c) in self.shapes(nu, nc).items():
            out[name] = np.array(rec[off[name]: off[name] + r * c]).reshape(r, c, order="F")
        return out


@dataclass(frozen=True)
class GainLayout:
    """``StageFactor::{ff, fb, fth}`` (riccati-kernel.hpp:76-78); fb / fth ROW-major."""
    nx: int
    nu: int
    nc: int
    nth: int

    @property
    def nrows(self) -> int:
        return self.nu + self.nc + 2 * self.nx

    @property
    def offsets(self):
        o = 0
        out = {"ff": o}
        o = _r2(o + self.nrows)
        out["fb"] = o
        o = _r2(o + self.nrows * self.nx)
        out["fth"] = o
        o = _r2(o + self.nrows * self.nth)
        out["stride"] = o
        return out

    @property
    def stride(self) -> int:
        return self.offsets["stride"]

    def unpack(self, rec: np.ndarray, nu=None, nc=None):
        """-> (ff [nr], fb [nr, nx], fth [nr, nth]) with nr = nu_t + nc_t + 2 nx."""
        nu = self.nu if nu is None else nu
        nc = self.nc if nc is None else nc
        nr = nu + nc + 2 * self.nx
        off = self.offsets
        ff = np.array(rec[off["ff"]: off["ff"] + nr])
        fb = np.array(rec[off["fb"]: off["fb"] + nr * self.nx]).reshape(nr, self.nx)
        fth = np.array(rec[off["fth"]: off["fth"] + nr * self.nth]).reshape(nr, self.nth)
        return ff, fb, fth


@dataclass(frozen=True)
class ValueLayout:
    """``StageFactor::value_t`` (riccati-kernel.hpp:39-60), column-major blocks."""
    nx: int
    nth: int

    def shapes(self):
        nx, nth = self.nx, self.nth
        return {"Pmat": (nx, nx), "pvec": (nx, 1), "Vxx": (nx, nx), "vx": (nx, 1),
                "Vxt": (nx, nth), "Vtt": (nth, nth), "vt": (nth, 1)}

    @property
    def offsets(self):
        o = 0
        out = {}
        for name, (r, c) in self.shapes().items():
            out[name] = o
            o = _r2(o + r * c)
        out["stride"] = o
        return out

    @property
    def stride(self) -> int:
        return self.offsets["stride"]

    def unpack(self, rec: np.ndarray) -> dict:
        off = self.offsets
        return {name: np.array(rec[off[name]: off[name] + r * c]).reshape(r, c, order="F")
                for name, (r, c) in self.shapes().items()}


def traj_sizes(nx: int, nu: int, nc: int, nc0: int, T: int):
    """Per-problem lengths of the packed trajectory arrays (xs, us, vs, lbdas).

    xs (T+1)*nx | us (T+1)*nu | vs (T+1)*nc | lbdas nc0 + T*nx  — ``lbdas[0]`` has nc0
    entries (utils.hpp:150), every later co-state nx.
    """
    return (T + 1) * nx, (T + 1) * nu, (T + 1) * nc, nc0 + T * nx
